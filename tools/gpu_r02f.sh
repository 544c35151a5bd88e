#!/bin/bash
# round 2, capture f (2 GPUs): bench under torchrun after the gather fix; N = 1 and N = 2 strong-block hashes must agree
set -x
mkdir -p gpurun_out
timeout 900 python bench.py --steps 3 --warmup 3 --circuits off --no-cpu-baseline --strong-rows 0 > gpurun_out/r02f_bench_quick.json 2> gpurun_out/r02f_bench_quick.err; echo rc=$?
timeout 900 python bench.py --steps 3 --warmup 3 --circuit-rows 2048 --strong-rows 4096 --cpu-gates-per-core 16 > gpurun_out/r02f_bench_n1.json 2> gpurun_out/r02f_bench_n1.err; echo rc=$?; tail -3 gpurun_out/r02f_bench_n1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 --strong-rows 4096 > gpurun_out/r02f_bench_n2.json 2> gpurun_out/r02f_bench_n2.err; echo rc=$?; tail -5 gpurun_out/r02f_bench_n2.err
python - <<'PY'
import json
a=json.load(open('gpurun_out/r02f_bench_n1.json')); b=json.load(open('gpurun_out/r02f_bench_n2.json'))
print("N1", a['value'], a['strong']['rows_per_s'], a['strong']['output_sha256'])
print("N2", b['value'], b['strong']['rows_per_s'], b['strong']['output_sha256'])
print("HASH_EQUAL", a['strong']['output_sha256']==b['strong']['output_sha256'])
PY
