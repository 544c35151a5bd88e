#!/bin/bash
# round 2, capture b (2 GPUs): full GPU suite on the reworked backend (chunk pipeline, NCCL key broadcast, shapes, gRPC),
# shape sweep, bench
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r02b_gpus.txt
python -m pytest tests -m gpu -x -q --durations=12 > gpurun_out/r02b_gpu_tests.log 2>&1
tail -15 gpurun_out/r02b_gpu_tests.log
python tools/shape_sweep.py > gpurun_out/r02b_shape_sweep.jsonl 2> gpurun_out/r02b_shape_sweep.err
tail -3 gpurun_out/r02b_shape_sweep.jsonl; tail -3 gpurun_out/r02b_shape_sweep.err
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02b_bench.json 2> gpurun_out/r02b_bench.err
head -c 400 gpurun_out/r02b_bench.json; tail -5 gpurun_out/r02b_bench.err
