"""GPU tests of the drop-in seam: IWorkerGroup::schedule_task(MapTask) through herdb_map_async, with
herdsman's on-disk layout -- <key_dir>/<session>/1.key and <storage_dir>/<session>/<df>/<partition>
(key_service.cpp:28-31, storage_service.cpp:19-28) -- checked against the oracle's per-row interpreter."""
import os
import uuid

import numpy as np
import pytest

from herdsman_b200 import api, circuit as cm
import clientsim as cs

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def world(tmp_path_factory, test_key, oracle_lib):
    root = tmp_path_factory.mktemp("herd")
    key_dir, storage = root / "keys", root / "storage"
    session = str(uuid.uuid4())
    (key_dir / session).mkdir(parents=True)
    (key_dir / session / "1.key").write_bytes(test_key["blob"].tobytes())        # SchemaType::BINFHE == 1
    wg = api.CudaWorkerGroup(str(storage), str(key_dir), devices=(0,))
    yield {"wg": wg, "session": session, "storage": storage, "oracle": oracle_lib.Oracle(test_key["blob"]), **test_key}
    wg.close()


def _frame(w, rows, parts):
    """create_data_frame + append_to_data_frame: deal rows to partition files 0..P-1 in order"""
    df = str(uuid.uuid4())
    d = w["storage"] / w["session"] / df
    d.mkdir(parents=True)
    start = 0
    for k in range(parts):
        cnt = api.partition_rows(rows.shape[0], parts, k)
        api.partition_write(d / str(k), rows[start:start + cnt])
        start += cnt
    return df


def _task(w, c, in_df, out_df, part, rows):
    return api.MapTask(w["session"], api.InputDataFramePtr(api.DataFramePtr(in_df, part), rows),
                       api.DataFramePtr(out_df, part), api.CryptoKeyPtr(1), c)


def test_map_tasks_over_partitions_match_oracle(world):
    w = world
    c = cm.adder_circuit(8)
    n_rows, parts = 11, 3                                        # ragged: 4 + 4 + 3
    rng = np.random.default_rng(0)
    x = rng.integers(0, 256, n_rows); y = rng.integers(0, 256, n_rows)
    bits = np.concatenate([cm.pack_words(x, 8), cm.pack_words(y, 8)], axis=1)
    rows = cs.encrypt_bits(w["n"], w["sk"], bits, seed=40)
    in_df = _frame(w, rows, parts)
    out_df = str(uuid.uuid4())
    (w["storage"] / w["session"] / out_df).mkdir(parents=True)   # ExecutionService creates the output frame dir
    handles, seen = [], []
    for k in range(parts):
        h = w["wg"].schedule_task(_task(w, c, in_df, out_df, k, api.partition_rows(n_rows, parts, k)))
        h.set_completion_callback(lambda st, k=k: seen.append((k, st)))           # installed AFTER scheduling (executor.cpp:86-93)
        handles.append(h)
    assert w["wg"].concurrent_workers() >= 1
    for h in handles:
        assert h.wait(300) == api.TaskHandle.Status.COMPLETED and h.completed()
    assert sorted(seen) == [(k, api.TaskHandle.Status.COMPLETED) for k in range(parts)]
    got = np.concatenate([api.partition_read(w["storage"] / w["session"] / out_df / str(k), api.partition_rows(n_rows, parts, k), 8, w["n"])
                          for k in range(parts)])
    assert not list((w["storage"] / w["session"] / out_df).glob("*.tmp"))
    assert np.array_equal(cm.unpack_words(cs.decrypt_bits(w["n"], w["sk"], got)), (x + y) % 256)
    assert np.array_equal(got, w["oracle"].circuit_rows(c.to_c(), rows, 8))       # bit-exact with the per-row interpreter


def test_chained_map_stages(world):
    """config 5 shape: add -> compare -> select as three map tasks over intermediate frames on disk"""
    w = world
    n_rows = 5
    rng = np.random.default_rng(1)
    a = rng.integers(0, 256, n_rows); b = rng.integers(0, 256, n_rows); t = rng.integers(0, 256, n_rows)
    D = cm.DataType
    # stage 1: (a, b, t) -> (a+b, t)
    s1 = cm.Circuit([[D.UINT8, D.UINT8, D.UINT8]], [("s", D.UINT8), ("t", D.UINT8)])
    s1.output_word(0, cm._ripple_add(s1, s1.input_word(0, 0), s1.input_word(0, 1), "wire")); s1.output_word(1, s1.input_word(0, 2))
    # stage 2: (s, t) -> (s < t, s, t)
    s2 = cm.Circuit([[D.UINT8, D.UINT8]], [("lt", D.BIT), ("s", D.UINT8), ("t", D.UINT8)])
    xs, ys = s2.input_word(0, 0), s2.input_word(0, 1)
    s2.output_bit(0, 0, cm._less_than(s2, xs, ys, "wire")); s2.output_word(1, xs); s2.output_word(2, ys)
    s3 = cm.select_circuit(8)
    bits = np.concatenate([cm.pack_words(a, 8), cm.pack_words(b, 8), cm.pack_words(t, 8)], axis=1)
    rows = cs.encrypt_bits(w["n"], w["sk"], bits, seed=41)
    df = _frame(w, rows, 1)
    ref = rows
    for c in (s1, s2, s3):
        out_df = str(uuid.uuid4()); (w["storage"] / w["session"] / out_df).mkdir(parents=True)
        assert w["wg"].schedule_task(_task(w, c, df, out_df, 0, n_rows)).wait(300) == api.TaskHandle.Status.COMPLETED
        ref = w["oracle"].circuit_rows(c.to_c(), ref, c.out_bits)
        df = out_df
    got = api.partition_read(w["storage"] / w["session"] / df / "0", n_rows, 8, w["n"])
    assert np.array_equal(got, ref)
    s = (a + b) % 256
    assert np.array_equal(cm.unpack_words(cs.decrypt_bits(w["n"], w["sk"], got)), np.where(s < t, s, t))


def test_empty_partition_and_error_statuses(world):
    w = world
    c = cm.single_gate_circuit()
    in_df = _frame(w, np.zeros((0, 2, w["n"] + 1), np.uint16), 1)
    out_df = str(uuid.uuid4()); (w["storage"] / w["session"] / out_df).mkdir(parents=True)
    assert w["wg"].schedule_task(_task(w, c, in_df, out_df, 0, 0)).wait(120) == api.TaskHandle.Status.COMPLETED
    assert (w["storage"] / w["session"] / out_df / "0").read_bytes() == b""
    # missing input partition, wrong row_count, wrong row payload size, unknown session key => ERROR (job fails, executor.cpp:168-178)
    assert w["wg"].schedule_task(_task(w, c, in_df, out_df, 7, 1)).wait(120) == api.TaskHandle.Status.ERROR
    assert w["wg"].schedule_task(_task(w, c, in_df, out_df, 0, 3)).wait(120) == api.TaskHandle.Status.ERROR
    add = cm.adder_circuit(8)
    rows = cs.encrypt_bits(w["n"], w["sk"], np.zeros((2, 2), np.uint8), seed=42)
    df2 = _frame(w, rows, 1)
    assert w["wg"].schedule_task(_task(w, add, df2, out_df, 0, 2)).wait(120) == api.TaskHandle.Status.ERROR
    bad = _task(w, c, df2, out_df, 0, 2); bad.session_uuid = str(uuid.uuid4())
    assert w["wg"].schedule_task(bad).wait(120) == api.TaskHandle.Status.ERROR
    # malformed circuit is refused at schedule time and reported through the handle
    broken = cm.Circuit([[cm.DataType.BIT]], [("o", cm.DataType.BIT)]); broken.output_bit(0, 0, broken.input(0, 5, 0))
    assert w["wg"].schedule_task(_task(w, broken, df2, out_df, 0, 2)).wait(120) == api.TaskHandle.Status.ERROR


def _oracle_tree_fold(orc, c, rows):
    """the backend's documented folding order (include/herdb.h, herdb_reduce_async) restated with the oracle:
    pairs (0,1),(2,3),... per round, an odd last row carried to the end of the next list"""
    cur = rows
    while cur.shape[0] > 1:
        m = cur.shape[0]
        pairs = cur[:m - (m & 1)].reshape(m // 2, 2 * cur.shape[1], cur.shape[2])
        nxt = orc.circuit_rows(c.to_c(), pairs, c.out_bits)
        cur = np.concatenate([nxt, cur[m - 1:]]) if m & 1 else nxt
    return cur


def test_reduce_task_over_partitions_matches_oracle(world):
    """IWorkerGroup::schedule_task(ReduceTask): rows of several input partitions folded to one output row"""
    w = world
    c = cm.reduce_add_circuit(8)
    n_rows, parts = 7, 3                                         # 3 + 2 + 2 rows, odd counts in the tree
    vals = np.random.default_rng(5).integers(0, 256, n_rows)
    rows = cs.encrypt_bits(w["n"], w["sk"], cm.pack_words(vals, 8), seed=43)
    in_df = _frame(w, rows, parts)
    out_df = str(uuid.uuid4()); (w["storage"] / w["session"] / out_df).mkdir(parents=True)
    ptrs = [api.InputDataFramePtr(api.DataFramePtr(in_df, k), api.partition_rows(n_rows, parts, k)) for k in range(parts)]
    task = api.ReduceTask(w["session"], ptrs, api.DataFramePtr(out_df, 0), api.CryptoKeyPtr(1), c)
    assert w["wg"].schedule_task(task).wait(300) == api.TaskHandle.Status.COMPLETED
    got = api.partition_read(w["storage"] / w["session"] / out_df / "0", 1, 8, w["n"])
    assert cm.unpack_words(cs.decrypt_bits(w["n"], w["sk"], got))[0] == vals.sum() % 256
    assert np.array_equal(got, _oracle_tree_fold(w["oracle"], c, rows))           # bit-exact, same folding order
    # a single input row passes through unchanged; a map-shaped circuit is refused
    one = _frame(w, rows[:1], 1)
    t1 = api.ReduceTask(w["session"], [api.InputDataFramePtr(api.DataFramePtr(one, 0), 1)], api.DataFramePtr(out_df, 1), api.CryptoKeyPtr(1), c)
    assert w["wg"].schedule_task(t1).wait(120) == api.TaskHandle.Status.COMPLETED
    assert np.array_equal(api.partition_read(w["storage"] / w["session"] / out_df / "1", 1, 8, w["n"]), rows[:1])
    bad = api.ReduceTask(w["session"], ptrs, api.DataFramePtr(out_df, 2), api.CryptoKeyPtr(1), cm.adder_circuit(8))
    assert w["wg"].schedule_task(bad).wait(120) == api.TaskHandle.Status.ERROR
    none = api.ReduceTask(w["session"], [], api.DataFramePtr(out_df, 3), api.CryptoKeyPtr(1), c)
    assert w["wg"].schedule_task(none).wait(120) == api.TaskHandle.Status.ERROR


def test_reduce_rows_max(world):
    w = world
    c = cm.reduce_max_circuit(8)
    vals = np.random.default_rng(6).integers(0, 256, 6)
    rows = cs.encrypt_bits(w["n"], w["sk"], cm.pack_words(vals, 8), seed=44)
    got = w["wg"].backend.reduce_rows(w["session"], c, rows)
    assert cm.unpack_words(cs.decrypt_bits(w["n"], w["sk"], got))[0] == vals.max()
    assert np.array_equal(got, _oracle_tree_fold(w["oracle"], c, rows))


def test_pipeline_rows_keep_intermediates_on_device(world):
    """herdb_eval_pipeline_rows: add -> compare -> select with the intermediate rows resident in HBM gives the
    same ciphertexts as the three stages evaluated one by one (and as the oracle)"""
    w = world
    n_rows = 6
    rng = np.random.default_rng(8)
    a = rng.integers(0, 256, n_rows); b = rng.integers(0, 256, n_rows); t = rng.integers(0, 256, n_rows)
    s1, s2, s3 = cm.pipeline_stages(8)
    bits = np.concatenate([cm.pack_words(a, 8), cm.pack_words(b, 8), cm.pack_words(t, 8)], axis=1)
    rows = cs.encrypt_bits(w["n"], w["sk"], bits, seed=45)
    be = w["wg"].backend
    be.load_key(w["session"])
    got = be.eval_pipeline_rows(w["session"], [s1, s2, s3], rows)
    ref = rows
    for c in (s1, s2, s3):
        ref = w["oracle"].circuit_rows(c.to_c(), ref, c.out_bits)
    assert np.array_equal(got, ref)
    s = (a + b) % 256
    assert np.array_equal(cm.unpack_words(cs.decrypt_bits(w["n"], w["sk"], got)), np.where(s < t, s, t))
    with pytest.raises(api.HerdbError):                          # stage 2 does not take stage 1's row
        be.eval_pipeline_rows(w["session"], [s1, s3], rows)
