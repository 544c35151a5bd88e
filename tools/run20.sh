# capture p: + key words rotated against register bank conflicts (final code). GPU tests, smoke, bench (both arms), launch list, ncu full of the two hot kernels
set -x
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3) > gpurun_out/r01p_tests.log 2>&1; cat gpurun_out/r01p_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r01p_smoke.log 2>&1; tail -1 gpurun_out/r01p_smoke.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r01p_bench.json 2> gpurun_out/r01p_bench.err; echo rc=$?; tail -3 gpurun_out/r01p_bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r01p_bench_reference_arm.json 2> gpurun_out/r01p_bench_reference.err; echo rc=$?
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01p_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r01p_ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"blind_rotate" --launch-skip 1 -c 1 -f -o gpurun_out/r01p_full_big python tools/prof_gate.py 33152 1 > gpurun_out/r01p_ncu_full_big.log 2>&1; tail -2 gpurun_out/r01p_ncu_full_big.log
timeout 900 python tools/bench_circuits.py --oracle-rows 2 --rows-lt32 4096 --rows-mul32 512 > gpurun_out/r01p_circuits.jsonl 2> gpurun_out/r01p_circuits.err; echo rc=$?
