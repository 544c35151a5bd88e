"""Multi-device identity: one process that owns G GPUs cuts a task's rows into
per-device slices; rows and the output partition FILE must be byte-identical to the one-GPU run, and the first
rows bit-exact with the oracle.  Keys reach devices 1..G-1 by the in-library NCCL broadcast (herdb_load_key).
Skipped on a one-GPU box; run by hand with `gpurun --gpus 2|8` (logs under profiles/)."""
import hashlib
import uuid

import numpy as np
import pytest

from herdsman_b200 import api, circuit as cm
import clientsim as cs

pytestmark = pytest.mark.gpu


def _device_count():
    import torch
    return torch.cuda.device_count()


def _world(tmp_path, key, devices):
    key_dir, storage = tmp_path / "keys", tmp_path / "storage"
    session = "5e551011-0000-4000-8000-%012d" % len(devices)
    (key_dir / session).mkdir(parents=True, exist_ok=True)
    (key_dir / session / "1.key").write_bytes(key["blob"].tobytes())
    wg = api.CudaWorkerGroup(str(storage), str(key_dir), devices=devices)
    return wg, session, storage


def _run_map(wg, session, storage, c, rows, n):
    in_df, out_df = str(uuid.uuid4()), str(uuid.uuid4())
    (storage / session / in_df).mkdir(parents=True); (storage / session / out_df).mkdir(parents=True)
    api.partition_write(storage / session / in_df / "0", rows)
    t = api.MapTask(session, api.InputDataFramePtr(api.DataFramePtr(in_df, 0), rows.shape[0]), api.DataFramePtr(out_df, 0), api.CryptoKeyPtr(1), c)
    assert wg.schedule_task(t).wait(600) == api.TaskHandle.Status.COMPLETED
    raw = (storage / session / out_df / "0").read_bytes()
    return raw, api.partition_read(storage / session / out_df / "0", rows.shape[0], c.out_bits, n)


@pytest.mark.parametrize("G", [2, 4, 8])
def test_rows_and_partition_file_identical_on_G_devices(tmp_path, std128_key, oracle_lib, G):
    if _device_count() < G:
        pytest.skip("needs %d GPUs" % G)
    k = std128_key
    c = cm.compare_circuit(32)
    rows_n = 37 * G + 5                                   # ragged slices
    rng = np.random.default_rng(3)
    x = rng.integers(0, 1 << 32, rows_n, dtype=np.uint64); y = rng.integers(0, 1 << 32, rows_n, dtype=np.uint64)
    y[:3] = x[:3]
    rows = cs.encrypt_bits(k["n"], k["sk"], np.concatenate([cm.pack_words(x, 32), cm.pack_words(y, 32)], axis=1), seed=3)
    wg1, s1, st1 = _world(tmp_path / "one", k, (0,))
    try:
        raw1, out1 = _run_map(wg1, s1, st1, c, rows, k["n"])
        host1 = wg1.backend.eval_circuit_rows(s1, c, rows)
    finally:
        wg1.close()
    wgG, sG, stG = _world(tmp_path / "many", k, tuple(range(G)))
    try:
        rawG, outG = _run_map(wgG, sG, stG, c, rows, k["n"])
        hostG = wgG.backend.eval_circuit_rows(sG, c, rows)
        stats = [wgG.backend.stats(d) for d in range(G)]
    finally:
        wgG.close()
    assert all(s["gates_bootstrapped"] > 0 for s in stats), "every device must have evaluated its slice"
    assert hashlib.sha256(raw1).hexdigest() == hashlib.sha256(rawG).hexdigest()          # output partition file, byte for byte
    assert np.array_equal(out1, outG) and np.array_equal(host1, hostG) and np.array_equal(host1, out1)
    dec = cs.decrypt_bits(k["n"], k["sk"], outG).reshape(rows_n)
    assert np.array_equal(dec, (x < y).astype(dec.dtype))
    orc = oracle_lib.Oracle(k["blob"])
    sample = np.r_[0:4, rows_n - 4:rows_n]                # first slice and last slice
    assert np.array_equal(outG[sample], orc.circuit_rows(c.to_c(), rows[sample], 1))


def test_pipeline_and_reduce_identical_on_two_devices(tmp_path, test_key):
    if _device_count() < 2:
        pytest.skip("needs 2 GPUs")
    k = test_key
    rng = np.random.default_rng(9)
    rows_n = 23
    a = rng.integers(0, 256, rows_n); b = rng.integers(0, 256, rows_n); t = rng.integers(0, 256, rows_n)
    rows = cs.encrypt_bits(k["n"], k["sk"], np.concatenate([cm.pack_words(v, 8) for v in (a, b, t)], axis=1), seed=9)
    red = cs.encrypt_bits(k["n"], k["sk"], cm.pack_words(a, 8), seed=10)
    outs = []
    for devs in [(0,), (0, 1), (1, 0)]:
        be = api.Backend(devices=devs)
        try:
            be.load_key_mem("s", k["blob"])
            outs.append((be.eval_pipeline_rows("s", cm.pipeline_stages(8), rows), be.reduce_rows("s", cm.reduce_add_circuit(8), red)))
        finally:
            be.close()
    for p, r in outs[1:]:
        assert np.array_equal(p, outs[0][0]) and np.array_equal(r, outs[0][1])
    s = (a + b) % 256
    assert np.array_equal(cm.unpack_words(cs.decrypt_bits(k["n"], k["sk"], outs[0][0])), np.where(s < t, s, t))
