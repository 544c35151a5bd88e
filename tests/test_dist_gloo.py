"""N > 1 host logic on CPU: world_size-2 gloo run of key-blob broadcast + row sharding.  The hot path
has no data-path collective (rows are independent); each rank evaluates its slice and the
concatenation in rank order must equal the single-process result.  (On CPU the per-slice evaluation
is done by the oracle: this test covers the sharding/broadcast plumbing, not the kernels.)"""
import hashlib
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from herdsman_b200 import api, circuit as cm, sharding
    import clientsim as cs
    from oracle import pyoracle as po
    n = api.N_TEST
    blob = sk = None
    if rank == 0:
        sk, skr, blob = cs.keygen(n)
    t = sharding.broadcast_blob(blob, dist)                       # key blob: one broadcast, then no collective
    blob = t.numpy()
    digest = hashlib.sha256(blob.tobytes()).hexdigest()
    sk_t = torch.from_numpy(sk.copy()) if rank == 0 else torch.empty(n, dtype=torch.int8)
    dist.broadcast(sk_t, src=0); sk = sk_t.numpy()
    c = cm.single_gate_circuit(cm.Operation.NAND)
    n_rows = 7                                                     # ragged: 4 + 3
    bits = np.random.default_rng(0).integers(0, 2, (n_rows, 2)).astype(np.uint8)
    rows = cs.encrypt_bits(n, sk, bits, seed=5)           # same on every rank (deterministic generator)
    lo, hi = sharding.row_slice(n_rows, world, rank)
    orc = po.Oracle(blob)
    local = orc.circuit_rows(c.to_c(), rows[lo:hi], 1)
    full = sharding.gather_rows(local, n_rows, dist)
    if rank == 0:
        single = orc.circuit_rows(c.to_c(), rows, 1)
        q.put((digest, bool(np.array_equal(full, single)), (lo, hi)))
    else:
        q.put((digest, None, (lo, hi)))
    dist.destroy_process_group()


def test_two_rank_gloo_sharding_and_key_broadcast():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert res[0][0] == res[1][0]                                  # identical key blob on both ranks
    assert any(r[1] is True for r in res)                          # sharded == unsharded, byte for byte
    assert sorted(r[2] for r in res) == [(0, 4), (4, 7)]
