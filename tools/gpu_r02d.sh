#!/bin/bash
# round 2, capture d (1 GPU): two kernel streams per lane (chunk c and c+1 in flight together); full GPU suite, bench, task seam
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/r02d_gpu_tests.log 2>&1; tail -12 gpurun_out/r02d_gpu_tests.log
timeout 1200 python bench.py --steps 10 --warmup 3 > gpurun_out/r02d_bench.json 2> gpurun_out/r02d_bench.err; echo rc=$?; tail -3 gpurun_out/r02d_bench.err
for cfg in pipeline lt32; do
  timeout 900 python tools/bench_tasks.py --config $cfg --rows 16384 --partitions 4 >> gpurun_out/r02d_tasks.jsonl 2>> gpurun_out/r02d_tasks.err; echo rc=$?
done
timeout 900 python tools/bench_tasks.py --config add8 --rows 1024 --partitions 8 >> gpurun_out/r02d_tasks.jsonl 2>> gpurun_out/r02d_tasks.err; echo rc=$?
timeout 900 python tools/bench_tasks.py --config add8 --rows 1024 --partitions 1 >> gpurun_out/r02d_tasks.jsonl 2>> gpurun_out/r02d_tasks.err; echo rc=$?
timeout 900 python tools/bench_tasks.py --config add8 --rows 256 --partitions 2 >> gpurun_out/r02d_tasks.jsonl 2>> gpurun_out/r02d_tasks.err; echo rc=$?
tail -3 gpurun_out/r02d_tasks.err
