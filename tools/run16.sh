# capture m, part 2: ncu --set full of the MIDDLE piece (54 grid rounds) of a 33,152-gate batch, circuit configurations
set -x
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"blind_rotate" --launch-skip 1 -c 1 -f -o gpurun_out/r01m_full_big python tools/prof_gate.py 33152 1 > gpurun_out/r01m_ncu_full_big.log 2>&1; tail -2 gpurun_out/r01m_ncu_full_big.log
timeout 900 python tools/bench_circuits.py --oracle-rows 2 --rows-lt32 4096 --rows-mul32 512 > gpurun_out/r01m_circuits.jsonl 2> gpurun_out/r01m_circuits.err; echo rc=$?
timeout 600 python tools/bench_circuits.py --configs sweep --sweep-max 65536 > gpurun_out/r01m_sweep.jsonl 2> gpurun_out/r01m_sweep.err; echo rc=$?
