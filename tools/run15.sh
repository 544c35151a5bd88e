# capture m: FP64-fold kernel. GPU tests, smoke, bench (both arms), launch list, ncu full of the two hot kernels
set -x
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3) > gpurun_out/r01m_tests.log 2>&1; cat gpurun_out/r01m_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r01m_smoke.log 2>&1; tail -1 gpurun_out/r01m_smoke.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r01m_bench.json 2> gpurun_out/r01m_bench.err; echo rc=$?; tail -3 gpurun_out/r01m_bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r01m_bench_reference_arm.json 2> gpurun_out/r01m_bench_reference.err; echo rc=$?
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01m_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r01m_ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"blind_rotate|keyswitch" -c 2 -f -o gpurun_out/r01m_full python tools/prof_gate.py 33152 1 > gpurun_out/r01m_ncu_full.log 2>&1; tail -2 gpurun_out/r01m_ncu_full.log
