for sk in 0 100 200 400 800; do echo "skew $sk"; HERDB_BR_SKEW=$sk python tools/prof_gate.py 33152 2 | tail -2; done
