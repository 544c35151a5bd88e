# A/B: fold62 on the IMAD pipe (foldint) against the FP64-pipe quotient (foldf64b); same box, same clocks
set -x
nvidia-smi -L
for v in p rot2 p rot2; do
  cp tools/libherdb_$v.so herdsman_b200/libherdb.so
  echo "== $v"; timeout 300 python tools/prof_gate.py 33152 3
done
python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -5
