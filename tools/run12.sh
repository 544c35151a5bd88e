set -x
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3) > gpurun_out/r01l_tests.log 2>&1; cat gpurun_out/r01l_tests.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r01l_bench.json 2> gpurun_out/r01l_bench.err; echo rc=$?; tail -3 gpurun_out/r01l_bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r01l_bench_reference.json 2> gpurun_out/r01l_bench_reference.err; echo rc=$?
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01l_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r01l_ncu_launches.log 2>&1
timeout 1200 python tools/bench_circuits.py --configs sweep --sweep-max 1048576 > gpurun_out/r01l_sweep.jsonl 2> gpurun_out/r01l_sweep.err; echo rc=$?
