set -x
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3) > gpurun_out/r01k_tests.log 2>&1; cat gpurun_out/r01k_tests.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r01k_bench.json 2> gpurun_out/r01k_bench.err; echo rc=$?; tail -3 gpurun_out/r01k_bench.err
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r01k_bench2.json 2> gpurun_out/r01k_bench2.err; echo rc=$?
