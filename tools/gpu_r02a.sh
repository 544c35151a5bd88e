#!/bin/bash
# round 2, capture a: new parity tests on the round-1 library, compute-sanitizer over smoke(), bench baseline
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r02a_gpus.txt; nproc >> gpurun_out/r02a_gpus.txt
python -m pytest tests -m gpu -q --durations=15 > gpurun_out/r02a_gpu_tests.log 2>&1
tail -5 gpurun_out/r02a_gpu_tests.log
for tool in memcheck racecheck synccheck initcheck; do
  timeout 1200 compute-sanitizer --tool $tool --log-file gpurun_out/r02a_sanitizer_$tool.log \
      python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/r02a_sanitizer_$tool.out 2>&1
  echo "$tool exit $?"; tail -3 gpurun_out/r02a_sanitizer_$tool.log
done
python bench.py --steps 5 --warmup 3 > gpurun_out/r02a_bench.json 2> gpurun_out/r02a_bench.err
cat gpurun_out/r02a_bench.json | head -c 600
