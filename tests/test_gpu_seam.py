"""GPU tests of the task seam's host runtime: the chunk pipeline (double-buffered row chunks, framed partition rows
used in place on the device), the reference's reduce-tree shapes, key-file revalidation / eviction, the small-job
CTA shapes and the two task threads.  Everything is compared with the oracle or with an unchunked run."""
import os
import sys
import threading
import uuid

import numpy as np
import pytest

from herdsman_b200 import api, circuit as cm
import clientsim as cs

sys.path.insert(0, os.path.dirname(__file__))
import reduce_trees as rt  # noqa: E402

pytestmark = pytest.mark.gpu


def _mk_world(root, key, **env):
    key_dir, storage = root / "keys", root / "storage"
    session = str(uuid.uuid4())
    (key_dir / session).mkdir(parents=True)
    (key_dir / session / "1.key").write_bytes(key["blob"].tobytes())
    old = {k: os.environ.get(k) for k in env}
    os.environ.update({k: str(v) for k, v in env.items()})
    try:
        wg = api.CudaWorkerGroup(str(storage), str(key_dir), devices=(0,))
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    return {"wg": wg, "session": session, "storage": storage, "key_dir": key_dir, **key}


@pytest.fixture(scope="module")
def world(tmp_path_factory, test_key, oracle_lib):
    w = _mk_world(tmp_path_factory.mktemp("seam"), test_key)
    w["oracle"] = oracle_lib.Oracle(test_key["blob"])
    yield w
    w["wg"].close()


def _new_df(w):
    df = str(uuid.uuid4())
    (w["storage"] / w["session"] / df).mkdir(parents=True)
    return df


def _map(w, c, rows, timeout=300, session=None):
    in_df, out_df = _new_df(w), _new_df(w)
    api.partition_write(w["storage"] / w["session"] / in_df / "0", rows)
    t = api.MapTask(session or w["session"], api.InputDataFramePtr(api.DataFramePtr(in_df, 0), rows.shape[0]),
                    api.DataFramePtr(out_df, 0), api.CryptoKeyPtr(1), c)
    st = w["wg"].schedule_task(t).wait(timeout)
    path = w["storage"] / w["session"] / out_df / "0"
    return st, (api.partition_read(path, rows.shape[0], c.out_bits, w["n"]) if st == api.TaskHandle.Status.COMPLETED else None)


def _adder_rows(w, rows, seed):
    rng = np.random.default_rng(seed)
    x = rng.integers(0, 256, rows); y = rng.integers(0, 256, rows)
    bits = np.concatenate([cm.pack_words(x, 8), cm.pack_words(y, 8)], axis=1)
    return x, y, cs.encrypt_bits(w["n"], w["sk"], bits, seed=seed)


def test_chunked_task_pipeline_matches_oracle(tmp_path, test_key, oracle_lib):
    """HERDB_MEM_BUDGET_MB=1 -> 13 adder rows per chunk: a 70-row partition runs as 6 double-buffered chunks (pread ->
    pinned -> H2D -> kernels on the framed rows -> framed output image -> D2H -> pwrite); 1 MiB of I/O budget cuts a
    host-buffer call the same way.  Output file and rows equal the oracle's, byte for byte."""
    w = _mk_world(tmp_path, test_key, HERDB_MEM_BUDGET_MB=1, HERDB_IO_CHUNK_MB=1)
    try:
        c = cm.adder_circuit(8)
        x, y, rows = _adder_rows(w, 70, 31)
        w["wg"].backend.reset_stats()
        st, got = _map(w, c, rows)
        assert st == api.TaskHandle.Status.COMPLETED
        stats = w["wg"].backend.stats()
        assert stats["blind_rotate_launches"] >= 4 * api.analyze(c)["levels"]            # really chunked
        ref = oracle_lib.Oracle(test_key["blob"]).circuit_rows(c.to_c(), rows, 8)
        assert np.array_equal(got, ref)
        assert np.array_equal(cm.unpack_words(cs.decrypt_bits(w["n"], w["sk"], got)), (x + y) % 256)
        assert stats["h2d_bytes"] >= rows.nbytes + 4 * 70 and stats["d2h_bytes"] >= got.nbytes + 4 * 70   # framed bytes moved as they are
        assert np.array_equal(w["wg"].backend.eval_circuit_rows(w["session"], c, rows), ref)  # pageable host rows, staged through pinned buffers
        import torch
        pin_in = torch.from_numpy(rows.copy()).pin_memory(); pin_out = torch.empty((70, 8, w["n"] + 1), dtype=torch.uint16).pin_memory()
        w["wg"].backend.eval_circuit_rows(w["session"], c, pin_in, out=pin_out)              # page-locked caller rows: DMA straight from / to them
        assert np.array_equal(pin_out.numpy(), ref)
        piped = w["wg"].backend.eval_pipeline_rows(w["session"], cm.pipeline_stages(8), np.concatenate([rows, rows[:, :8]], axis=1))
        assert piped.shape == (70, 8, w["n"] + 1)
        w["wg"].backend.unload_key(w["session"])          # evaluations pin the key only while they run
        assert w["wg"].backend.loaded_keys() == 0
    finally:
        w["wg"].close()


def test_two_tasks_in_flight_and_zero_output_columns(world):
    """concurrent_workers() == 2 == the backend's task threads: partitions scheduled back to back all complete, results
    are per-task correct"""
    w = world
    assert w["wg"].backend.task_threads() == 2 and w["wg"].concurrent_workers() == w["wg"].backend.task_window() == 8
    c = cm.adder_circuit(8)
    jobs = []
    for k in range(5):
        x, y, rows = _adder_rows(w, 3 + k, 50 + k)
        in_df, out_df = _new_df(w), _new_df(w)
        api.partition_write(w["storage"] / w["session"] / in_df / "0", rows)
        t = api.MapTask(w["session"], api.InputDataFramePtr(api.DataFramePtr(in_df, 0), rows.shape[0]), api.DataFramePtr(out_df, 0), api.CryptoKeyPtr(1), c)
        jobs.append((w["wg"].schedule_task(t), out_df, rows))
    for h, out_df, rows in jobs:
        assert h.wait(300) == api.TaskHandle.Status.COMPLETED
        got = api.partition_read(w["storage"] / w["session"] / out_df / "0", rows.shape[0], 8, w["n"])
        assert np.array_equal(got, w["oracle"].circuit_rows(c.to_c(), rows, 8))


def test_small_map_tasks_are_coalesced_and_fail_individually(world):
    """eight 6-row partitions of one frame scheduled at once (the executor's window is concurrent_workers() = 8): the task
    threads evaluate them as one or two batches instead of eight (far fewer launches), every task still writes its own
    partition, equal to the oracle's; a task whose partition is missing, queued in the middle of such a group, fails alone"""
    w = world
    be = w["wg"].backend
    c = cm.adder_circuit(8)
    levels = api.analyze(c)["levels"]
    parts, per = 8, 6
    x, y, rows = _adder_rows(w, parts * per, 123)
    in_df, out_df = _new_df(w), _new_df(w)
    for k in range(parts):
        if k != 3:                                                  # partition 3 does not exist
            api.partition_write(w["storage"] / w["session"] / in_df / str(k), rows[k * per:(k + 1) * per])
    be.wait_idle(); be.reset_stats()
    handles = [w["wg"].schedule_task(api.MapTask(w["session"], api.InputDataFramePtr(api.DataFramePtr(in_df, k), per),
                                                 api.DataFramePtr(out_df, k), api.CryptoKeyPtr(1), c)) for k in range(parts)]
    st = [h.wait(300) for h in handles]
    assert [s == api.TaskHandle.Status.COMPLETED for s in st] == [k != 3 for k in range(parts)]
    ref = w["oracle"].circuit_rows(c.to_c(), rows, 8)
    for k in range(parts):
        path = w["storage"] / w["session"] / out_df / str(k)
        if k == 3:
            assert not path.exists() and not path.with_name("3.tmp").exists()
        else:
            assert np.array_equal(api.partition_read(path, per, 8, w["n"]), ref[k * per:(k + 1) * per])
    # uncoalesced: 7 tasks x `levels` blind-rotation launches; a failed group is re-run task by task (the fallback), so
    # only a run without the bad task shows the saving
    be.wait_idle(); be.reset_stats()
    out2 = _new_df(w)
    api.partition_write(w["storage"] / w["session"] / in_df / "3", rows[3 * per:4 * per])
    handles = [w["wg"].schedule_task(api.MapTask(w["session"], api.InputDataFramePtr(api.DataFramePtr(in_df, k), per),
                                                 api.DataFramePtr(out2, k), api.CryptoKeyPtr(1), c)) for k in range(parts)]
    assert all(h.wait(300) == api.TaskHandle.Status.COMPLETED for h in handles)
    got = np.concatenate([api.partition_read(w["storage"] / w["session"] / out2 / str(k), per, 8, w["n"]) for k in range(parts)])
    assert np.array_equal(got, ref)
    assert be.stats()["blind_rotate_launches"] <= parts * levels // 2, "tasks were not coalesced"


@pytest.mark.parametrize("policy,parts,per_node", [("SEQUENCED", 3, 2), ("PARALLEL", 3, 2), ("PARALLEL_FULL", 5, 2), ("PARALLEL_FULL", 7, 3)])
def test_reference_reduce_trees(world, policy, parts, per_node):
    """the ReduceTask sequences herdsman issues for Policy::SEQUENCED / PARALLEL / PARALLEL_FULL
    (execution_service.cpp:560-689, restated in tests/reduce_trees.py), with the reference's own frame labels:
    hidden partitions labelled row_count = partitions hold one row each, the final output goes to partition 1"""
    w = world
    c = cm.reduce_add_circuit(8)
    n_rows = 2 * parts + 1
    vals = np.random.default_rng(parts).integers(0, 256, n_rows)
    rows = cs.encrypt_bits(w["n"], w["sk"], cm.pack_words(vals, 8), seed=90 + parts)
    in_df, out_df, hid_df = _new_df(w), _new_df(w), _new_df(w)
    part_rows, start = [], 0
    for k in range(parts):
        cnt = api.partition_rows(n_rows, parts, k)
        api.partition_write(w["storage"] / w["session"] / in_df / str(k), rows[start:start + cnt])
        part_rows.append(cnt); start += cnt
    for ins, out in rt.build(policy, parts, part_rows, in_df, out_df, hid_df, per_node):
        ptrs = [api.InputDataFramePtr(api.DataFramePtr(df, part), cnt) for df, cnt, part in ins]
        t = api.ReduceTask(w["session"], ptrs, api.DataFramePtr(out[0], out[2]), api.CryptoKeyPtr(1), c)
        assert w["wg"].schedule_task(t).wait(300) == api.TaskHandle.Status.COMPLETED, (policy, ins, out)
    got = api.partition_read(w["storage"] / w["session"] / out_df / "1", 1, 8, w["n"])
    assert cm.unpack_words(cs.decrypt_bits(w["n"], w["sk"], got))[0] == vals.sum() % 256


def test_key_file_is_revalidated_reloaded_and_evicted(tmp_path, test_key):
    """KeyService::add_key overwrites a session's key file, remove_key deletes it (key_service.cpp:19-60): the next
    task must evaluate with the key that is on disk, or fail -- never complete with the stale one.  Keys of idle
    sessions are evicted over HERDB_KEY_BUDGET_MB and come back from their file."""
    w = _mk_world(tmp_path, test_key, HERDB_KEY_BUDGET_MB=60)            # a test-set key is ~25 MiB on the device
    try:
        be = w["wg"].backend
        c = cm.single_gate_circuit(cm.Operation.NAND)
        bits = np.array([[0, 0], [0, 1], [1, 0], [1, 1]], np.uint8)
        rows = cs.encrypt_bits(w["n"], w["sk"], bits, seed=3)
        st, got = _map(w, c, rows)
        assert st == api.TaskHandle.Status.COMPLETED
        assert cs.decrypt_bits(w["n"], w["sk"], got).ravel().tolist() == [1, 1, 1, 0]
        # the session uploads a new key: same path, other key material
        sk2, _, blob2 = cs.keygen(w["n"], seed=0xB200)
        (w["key_dir"] / w["session"] / "1.key").write_bytes(blob2.tobytes())
        rows2 = cs.encrypt_bits(w["n"], sk2, bits, seed=4)
        st, got2 = _map(w, c, rows2)
        assert st == api.TaskHandle.Status.COMPLETED
        assert cs.decrypt_bits(w["n"], sk2, got2).ravel().tolist() == [1, 1, 1, 0]      # evaluated under the NEW key
        # two more sessions: the budget holds two keys, the least recently used one goes
        others = []
        for k in range(2):
            s = str(uuid.uuid4()); (w["key_dir"] / s).mkdir(); (w["storage"] / s).mkdir()
            (w["key_dir"] / s / "1.key").write_bytes(w["blob"].tobytes())
            wk = dict(w); wk["session"] = s
            st, g = _map(wk, c, rows)
            assert st == api.TaskHandle.Status.COMPLETED and cs.decrypt_bits(w["n"], w["sk"], g).ravel().tolist() == [1, 1, 1, 0]
            others.append(s)
        assert be.loaded_keys() == 2
        st, got3 = _map(w, c, rows2)                                                      # evicted session: reloaded from its file
        assert st == api.TaskHandle.Status.COMPLETED and np.array_equal(got3, got2)
        # remove_key: the file is gone => the task fails instead of using the cached copy
        os.remove(w["key_dir"] / w["session"] / "1.key")
        st, _ = _map(w, c, rows2)
        assert st == api.TaskHandle.Status.ERROR
        # CryptoKeyPtr.schema_type names the file: <schema_type>.key
        in_df, out_df = _new_df(w), _new_df(w)
        api.partition_write(w["storage"] / w["session"] / in_df / "0", rows)
        (w["key_dir"] / w["session"] / "7.key").write_bytes(w["blob"].tobytes())
        t = api.MapTask(w["session"], api.InputDataFramePtr(api.DataFramePtr(in_df, 0), 4), api.DataFramePtr(out_df, 0), api.CryptoKeyPtr(7), c)
        assert w["wg"].schedule_task(t).wait(120) == api.TaskHandle.Status.COMPLETED
        g7 = api.partition_read(w["storage"] / w["session"] / out_df / "0", 4, 1, w["n"])
        assert cs.decrypt_bits(w["n"], w["sk"], g7).ravel().tolist() == [1, 1, 1, 0]
    finally:
        w["wg"].close()


@pytest.mark.parametrize("batch", [1, 3, 150, 297, 600, 1000])
def test_small_job_cta_shapes_give_identical_ciphertexts(world, batch):
    """blind_rotate_kernel<2> / <1> (launch tails and small launches) against <4> and the oracle; the launch policy
    (body at 4, tail at 2 or 1) too"""
    import torch
    w = world
    be = w["wg"].backend
    be.load_key(w["session"])
    rng = np.random.default_rng(batch)
    a = rng.integers(0, 2, batch).astype(np.uint8); b = rng.integers(0, 2, batch).astype(np.uint8)
    ca = cs.encrypt_bits(w["n"], w["sk"], a, seed=batch, first=0); cb = cs.encrypt_bits(w["n"], w["sk"], b, seed=batch, first=1 << 20)
    d_a = torch.from_numpy(ca).cuda(); d_b = torch.from_numpy(cb).cuda()
    outs = {}
    for shape in (4, 2, 1, 0):
        d_out = torch.zeros_like(d_a)
        be.eval_gate_batch_dev(w["session"], "NAND", d_a, d_b, d_out, batch, shape=shape)
        outs[shape] = d_out.cpu().numpy()
    k = min(batch, 16)
    ref = w["oracle"].gate("NAND", ca[:k], cb[:k])
    for shape in (4, 2, 1, 0):
        assert np.array_equal(outs[shape], outs[4]), shape
        assert np.array_equal(outs[shape][:k], ref)
    assert np.array_equal(cs.decrypt_bits(w["n"], w["sk"], outs[0]), 1 - (a & b))


def test_small_job_shapes_std128(std128_key, oracle_lib):
    import torch
    k = std128_key
    be = api.Backend(devices=(0,))
    try:
        be.load_key_mem("s", k["blob"])
        rng = np.random.default_rng(4)
        B = 9
        a = rng.integers(0, 2, B).astype(np.uint8); b = rng.integers(0, 2, B).astype(np.uint8)
        ca = cs.encrypt_bits(k["n"], k["sk"], a, seed=4, first=0); cb = cs.encrypt_bits(k["n"], k["sk"], b, seed=4, first=1 << 20)
        ref = oracle_lib.Oracle(k["blob"]).gate("XOR_FAST", ca, cb)
        d_a = torch.from_numpy(ca).cuda(); d_b = torch.from_numpy(cb).cuda()
        for shape in (1, 2, 4):
            d_out = torch.zeros_like(d_a)
            be.eval_gate_batch_dev("s", "XOR_FAST", d_a, d_b, d_out, B, shape=shape)
            assert np.array_equal(d_out.cpu().numpy(), ref), shape
    finally:
        be.close()


def test_concurrent_host_callers_and_tasks_share_a_device(world):
    """file-less callers on several threads while tasks run: every evaluation owns a lane (its own staging buffers and
    gate tables), the device's kernel stream serialises them"""
    w = world
    be = w["wg"].backend
    c = cm.adder_circuit(8)
    x, y, rows = _adder_rows(w, 9, 77)
    want = w["oracle"].circuit_rows(c.to_c(), rows, 8)
    errs, res = [], []

    def host_caller():
        try:
            for _ in range(3):
                res.append(be.eval_circuit_rows(w["session"], c, rows))
        except Exception as e:      # noqa: BLE001
            errs.append(e)

    def task_caller():
        try:
            for _ in range(2):
                st, got = _map(w, c, rows)
                assert st == api.TaskHandle.Status.COMPLETED
                res.append(got)
        except Exception as e:      # noqa: BLE001
            errs.append(e)

    th = [threading.Thread(target=host_caller) for _ in range(2)] + [threading.Thread(target=task_caller) for _ in range(2)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errs and len(res) == 10
    for r in res:
        assert np.array_equal(r, want)
