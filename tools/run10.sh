set -x
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3) > gpurun_out/r01j_tests.log 2>&1; cat gpurun_out/r01j_tests.log
timeout 900 python tools/bench_circuits.py --oracle-rows 2 --rows-lt32 4096 --rows-mul32 512 > gpurun_out/r01j_circuits.jsonl 2> gpurun_out/r01j_circuits.err; echo rc=$?
timeout 300 python tools/probe_overhead.py > gpurun_out/probe_overhead2.txt 2>&1; cat gpurun_out/probe_overhead2.txt
