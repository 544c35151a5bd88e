#!/bin/bash
# round 2, capture j (1 GPU): task coalescing -- seam / task / adapter / gRPC tests, then the small-partition task benches
set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_seam.py tests/test_gpu_task.py tests/test_gpu_adapter.py tests/test_gpu_worker_grpc.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r02j_gpu_tests.log 2>&1; tail -6 gpurun_out/r02j_gpu_tests.log
for spec in "add8 1024 8" "add8 256 2" "add8 1024 1" "lt32 4096 8"; do
  set -- $spec
  timeout 600 python tools/bench_tasks.py --config $1 --rows $2 --partitions $3 >> gpurun_out/r02j_tasks.jsonl 2>> gpurun_out/r02j_tasks.err; echo rc=$?
done
HERDB_COALESCE_ROWS=0 timeout 600 python tools/bench_tasks.py --config add8 --rows 1024 --partitions 8 >> gpurun_out/r02j_tasks_nocoalesce.jsonl 2>> gpurun_out/r02j_tasks.err; echo rc=$?
tail -3 gpurun_out/r02j_tasks.err
python - <<'PY'
import json
for f in ["gpurun_out/r02j_tasks.jsonl","gpurun_out/r02j_tasks_nocoalesce.jsonl"]:
    for l in open(f):
        if l.startswith("{"):
            t=json.loads(l); print(f[-22:], t["config"],t["rows"],t["partitions"],"seam %.1f rows/s %.0f gates/s | host %.1f | ratio %.4f"%(t["seam"]["rows_per_s"],t["seam"]["gates_per_s"],t["host"]["rows_per_s"],t["seam_over_host"]), t["seam_rows_equal_host_rows"])
PY
