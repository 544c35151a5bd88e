#!/bin/bash
# round 2, capture g (8 GPUs, one process): multi-device identity tests (2 / 4 / 8 devices), in-library NCCL key broadcast timing,
# config 3 at its full 1M rows through the task seam.  Every result is written to gpurun_out/ as soon as it exists.
set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r02g_gpus.txt; free -g >> gpurun_out/r02g_gpus.txt; df -h /dev/shm /tmp >> gpurun_out/r02g_gpus.txt; nproc >> gpurun_out/r02g_gpus.txt
timeout 300 python -m pytest tests/test_gpu_multidevice.py -m gpu -x -q -rs > gpurun_out/r02g_multidevice_tests.log 2>&1; tail -5 gpurun_out/r02g_multidevice_tests.log
ROWS=${1:-1048576}
timeout 420 python tools/bench_tasks.py --big --config lt32 --rows $ROWS --partitions 8 --devices 8 --dir /dev/shm > gpurun_out/r02g_config3_full.jsonl 2> gpurun_out/r02g_config3_full.err; echo rc=$?; tail -3 gpurun_out/r02g_config3_full.err
cat gpurun_out/r02g_config3_full.jsonl | cut -c1-1500
