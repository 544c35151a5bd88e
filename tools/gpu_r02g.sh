#!/bin/bash
# round 2, capture g (8 GPUs, one process): multi-device identity tests, key broadcast timing, config 3 at its full 1M rows through the
# task seam, the pipeline config through the seam against the host-buffer path
set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r02g_gpus.txt; free -g >> gpurun_out/r02g_gpus.txt; df -h /dev/shm /tmp >> gpurun_out/r02g_gpus.txt; nproc >> gpurun_out/r02g_gpus.txt
timeout 900 python -m pytest tests/test_gpu_multidevice.py -m gpu -x -q -rs > gpurun_out/r02g_multidevice_tests.log 2>&1; tail -5 gpurun_out/r02g_multidevice_tests.log
MEM_GB=$(free -g | awk '/^Mem:/{print $7}')
ROWS=1048576; if [ "$MEM_GB" -lt 180 ]; then ROWS=262144; fi
timeout 1500 python tools/bench_tasks.py --big --config lt32 --rows $ROWS --partitions 8 --devices 8 --dir /dev/shm > gpurun_out/r02g_config3_full.jsonl 2> gpurun_out/r02g_config3_full.err; echo rc=$?; tail -3 gpurun_out/r02g_config3_full.err
timeout 900 python tools/bench_tasks.py --config pipeline --rows 131072 --partitions 8 --devices 8 --dir /dev/shm > gpurun_out/r02g_tasks_8dev.jsonl 2> gpurun_out/r02g_tasks_8dev.err; echo rc=$?; tail -3 gpurun_out/r02g_tasks_8dev.err
