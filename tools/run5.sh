set -x
tools/microbench9 > gpurun_out/r01e_mb9.txt 2>&1; cat gpurun_out/r01e_mb9.txt
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5) > gpurun_out/r01e_tests.log 2>&1; cat gpurun_out/r01e_tests.log
timeout 1200 python tools/bench_circuits.py --oracle-rows 2 --rows-lt32 4096 --rows-mul32 512 > gpurun_out/r01e_circuits.jsonl 2> gpurun_out/r01e_circuits.err; echo rc=$?; tail -2 gpurun_out/r01e_circuits.err
