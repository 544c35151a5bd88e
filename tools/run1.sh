set -x
nvidia-smi -L
(timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r01b_tests.log 2>&1; echo "tests exit $?" >> gpurun_out/r01b_tests.log)
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/r01b_bench.json 2> gpurun_out/r01b_bench.err
HERDB_BR_VARIANT=2 timeout 300 python tools/prof_gate.py 4736 2 > gpurun_out/r01b_var2.log 2>&1
timeout 300 python tools/prof_gate.py 4736 2 > gpurun_out/r01b_var4.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:blind_rotate -c 1 -f -o gpurun_out/r01b_br python tools/prof_gate.py 592 1 > gpurun_out/r01b_ncu.log 2>&1
for m in 4 5 6; do timeout 120 tools/microbench$m > gpurun_out/r01b_mb$m.txt 2>&1; done
tail -3 gpurun_out/r01b_tests.log; cat gpurun_out/r01b_bench.json; cat gpurun_out/r01b_var2.log gpurun_out/r01b_var4.log; cat gpurun_out/r01b_mb6.txt
