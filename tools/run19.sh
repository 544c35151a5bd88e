# profile of the experimental FP64-quotient forward pass 1 (not adopted): where does its time go?
set -x
cp herdsman_b200/libherdb.so /tmp/libherdb_product.so
cp tools/libherdb_p1f64c.so herdsman_b200/libherdb.so
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"blind_rotate" --launch-skip 1 -c 1 -f -o gpurun_out/r01p_p1f64c python tools/prof_gate.py 4736 1 > gpurun_out/r01p_p1f64c_ncu.log 2>&1; tail -2 gpurun_out/r01p_p1f64c_ncu.log
cp /tmp/libherdb_product.so herdsman_b200/libherdb.so
