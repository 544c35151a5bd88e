set -x
timeout 600 ncu --set full --clock-control none --import-source on -k regex:blind_rotate -c 1 -f -o gpurun_out/r01c_br python tools/prof_gate.py 592 1 > gpurun_out/r01c_ncu.log 2>&1
tail -3 gpurun_out/r01c_ncu.log
