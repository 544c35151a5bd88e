# quick GPU check: parity tests + kernel timing at one full wave x 8
set -x
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3)
timeout 300 python tools/prof_gate.py 4736 2
