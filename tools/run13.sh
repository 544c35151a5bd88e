set -x
for N in 2 4 8; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r01l_bench_${N}gpu.json 2> gpurun_out/r01l_bench_${N}gpu.err; echo rc=$?
done
timeout 900 python tools/bench_circuits.py --devices 8 --configs lt32,pipe8 --rows-lt32 65536 --rows-pipe 65536 > gpurun_out/r01l_circuits_8dev.jsonl 2> gpurun_out/r01l_circuits_8dev.err; echo rc=$?
