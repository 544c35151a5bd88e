#!/bin/bash
# A/B of library builds on one box: tools/ab.sh name1 name2 ...   ("base" = herdsman_b200/libherdb.so, x = libherdb_x.so)
# alternates the builds, three batches of 33,152 NAND gates each per turn; kernel times from the library's CUDA events
for rep in 1 2; do
  for v in "$@"; do
    lib=herdsman_b200/libherdb.so; [ "$v" != base ] && lib=herdsman_b200/libherdb_$v.so
    echo "== $v"
    HERDB_LIB=$PWD/$lib python tools/prof_gate.py 33152 3 | grep batch
  done
done
