#!/bin/bash
# round 2, capture e (2 GPUs): single kernel stream again; seam / multi-device / adapter / gRPC tests, bench under torchrun (strong block,
# hash equal at N = 1 and N = 2), the full-size mode of bench_tasks at a small size
set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_seam.py tests/test_gpu_multidevice.py tests/test_gpu_adapter.py tests/test_gpu_worker_grpc.py tests/test_gpu_task.py -m gpu -x -q > gpurun_out/r02e_gpu_tests.log 2>&1; tail -5 gpurun_out/r02e_gpu_tests.log
timeout 900 python bench.py --steps 3 --warmup 3 --circuit-rows 2048 --strong-rows 2048 --cpu-gates-per-core 16 > gpurun_out/r02e_bench_n1.json 2> gpurun_out/r02e_bench_n1.err; echo rc=$?; tail -3 gpurun_out/r02e_bench_n1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 --strong-rows 2048 > gpurun_out/r02e_bench_n2.json 2> gpurun_out/r02e_bench_n2.err; echo rc=$?; tail -5 gpurun_out/r02e_bench_n2.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 1 --warmup 1 > gpurun_out/r02e_bench_ref_n2.json 2> gpurun_out/r02e_bench_ref_n2.err; echo rc=$?
timeout 900 python tools/bench_tasks.py --big --config lt32 --rows 8192 --partitions 4 --devices 2 > gpurun_out/r02e_tasks_big.jsonl 2> gpurun_out/r02e_tasks_big.err; echo rc=$?; tail -3 gpurun_out/r02e_tasks_big.err
timeout 900 python tools/bench_tasks.py --config lt32 --rows 8192 --partitions 4 --devices 2 >> gpurun_out/r02e_tasks.jsonl 2>> gpurun_out/r02e_tasks.err; echo rc=$?
timeout 900 python tools/bench_tasks.py --config add8 --rows 1024 --partitions 8 >> gpurun_out/r02e_tasks.jsonl 2>> gpurun_out/r02e_tasks.err; echo rc=$?
df -h /tmp /dev/shm | tee gpurun_out/r02e_df.txt; free -g | tee -a gpurun_out/r02e_df.txt; nproc | tee -a gpurun_out/r02e_df.txt
