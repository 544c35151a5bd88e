set -x
nvidia-smi -L
python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -30
python - <<'PY'
import sys; sys.path.insert(0,'.')
import numpy as np, time
from herdsman_b200 import api
b = api.Backend()
print("int peaks", b.measure_int_peak())
sk, skr, blob = api.client_keygen(512)
t=time.time(); b.load_key_mem("s", blob); print("load key s", time.time()-t)
for batch in [592, 2368, 4736]:
    bits = np.random.default_rng(0).integers(0,2,batch).astype(np.uint8)
    ca = api.client_encrypt_bits(512, sk, bits, first=0); cb = api.client_encrypt_bits(512, sk, bits, first=1<<20)
    b.reset_stats(); t=time.time(); out = b.eval_gate_batch("s","NAND",ca,cb); dt=time.time()-t
    st=b.stats(); print(batch, "wall %.3fs"%dt, "gates/s %.0f"%(batch/dt), "br ms %.2f ks ms %.2f"%(st["blind_rotate_ms"], st["keyswitch_ms"]), "kernel gates/s %.0f"%(batch/((st["blind_rotate_ms"]+st["keyswitch_ms"])*1e-3)))
    assert (api.client_decrypt_bits(512, sk, out)==1-(bits&bits)).all()
PY
