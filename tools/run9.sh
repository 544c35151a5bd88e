set -x
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3) > gpurun_out/r01i_tests.log 2>&1; cat gpurun_out/r01i_tests.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r01i_bench.json 2> gpurun_out/r01i_bench.err; echo rc=$?; tail -3 gpurun_out/r01i_bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01i_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r01i_ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"blind_rotate|keyswitch" -c 2 -f -o gpurun_out/r01i_full python tools/prof_gate.py 33152 1 > gpurun_out/r01i_ncu_full.log 2>&1; tail -2 gpurun_out/r01i_ncu_full.log
timeout 900 python tools/bench_circuits.py --oracle-rows 2 --rows-lt32 4096 --rows-mul32 512 > gpurun_out/r01i_circuits.jsonl 2> gpurun_out/r01i_circuits.err; echo rc=$?
