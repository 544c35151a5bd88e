#!/bin/bash
# round 2, capture c (1 GPU): the full bench line (circuits, strong block, L2 roof), reference arm, task-seam bench,
# ncu launch list + full captures of the two hot kernels (each only after the same command exited 0 without ncu)
set -x
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02c_smoke.log 2>&1; tail -1 gpurun_out/r02c_smoke.log
timeout 1200 python bench.py --steps 10 --warmup 3 > gpurun_out/r02c_bench.json 2> gpurun_out/r02c_bench.err; echo rc=$?; tail -3 gpurun_out/r02c_bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02c_bench_reference_arm.json 2> gpurun_out/r02c_bench_reference.err; echo rc=$?
for cfg in pipeline lt32; do
  timeout 900 python tools/bench_tasks.py --config $cfg --rows 16384 --partitions 4 >> gpurun_out/r02c_tasks.jsonl 2>> gpurun_out/r02c_tasks.err; echo rc=$?
  timeout 900 python tools/bench_tasks.py --config $cfg --rows 16384 --partitions 4 --dir /dev/shm >> gpurun_out/r02c_tasks.jsonl 2>> gpurun_out/r02c_tasks.err; echo rc=$?
done
timeout 900 python tools/bench_tasks.py --config add8 --rows 1024 --partitions 8 >> gpurun_out/r02c_tasks.jsonl 2>> gpurun_out/r02c_tasks.err; echo rc=$?
tail -2 gpurun_out/r02c_tasks.err
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --circuits off > /dev/null 2> gpurun_out/r02c_prencu.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02c_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --circuits off > gpurun_out/r02c_ncu_launches.log 2>&1
timeout 300 python tools/prof_gate.py 33152 1 > gpurun_out/r02c_prof_gate.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"blind_rotate" --launch-skip 1 -c 1 -f -o /tmp/r02c_full_br python tools/prof_gate.py 33152 1 > gpurun_out/r02c_ncu_full_br.log 2>&1
ncu -i /tmp/r02c_full_br.ncu-rep --page details > gpurun_out/r02c_ncu_details_blind_rotate.txt 2>&1
ncu -i /tmp/r02c_full_br.ncu-rep --page raw --csv > gpurun_out/r02c_ncu_raw_blind_rotate.csv 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"keyswitch" --launch-skip 1 -c 1 -f -o /tmp/r02c_full_ks python tools/prof_gate.py 33152 1 > gpurun_out/r02c_ncu_full_ks.log 2>&1
ncu -i /tmp/r02c_full_ks.ncu-rep --page details > gpurun_out/r02c_ncu_details_keyswitch.txt 2>&1
ncu -i /tmp/r02c_full_ks.ncu-rep --page raw --csv > gpurun_out/r02c_ncu_raw_keyswitch.csv 2>&1
ls -la gpurun_out | tail -20
