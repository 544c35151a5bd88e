"""CPU tests of the host logic: circuit compiler, C-ABI surface, client stand-in, partition files."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from herdsman_b200 import api, circuit as cm
import clientsim as cs
from oracle import pyoracle as po

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "herdb.h")).read()
    declared = set(re.findall(r"^(?:int|void|size_t|uint64_t|const char\*)\s+(herdb_[a-z0-9_]+)\s*\(", hdr, re.M))
    assert declared == set(api.EXPORTS), declared ^ set(api.EXPORTS)
    L = api.lib()
    for name in declared:
        assert hasattr(L, name), name


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    rc = api.lib().herdb_create(C.byref(h), None, 0, b"", b"")
    assert rc == -2 and not h.value                      # HERDB_ERR_CUDA, loudly
    with pytest.raises(api.HerdbError):
        api.Backend()


def test_client_standin_matches_oracle_generator(test_key):
    sk2, skr2, blob2 = po.keygen(test_key["n"], 0x48455244)
    assert np.array_equal(test_key["blob"], blob2) and np.array_equal(test_key["sk"], sk2) and np.array_equal(test_key["skr"], skr2)
    bits = np.random.default_rng(0).integers(0, 2, 64).astype(np.uint8)
    c1 = cs.encrypt_bits(32, sk2, bits, seed=3, first=7)
    assert np.array_equal(c1, po.encrypt_bits(32, sk2, bits, seed=3, first=7))
    assert np.array_equal(cs.decrypt_bits(32, sk2, c1), bits)
    assert np.array_equal(po.decrypt_bits(32, sk2, c1), bits)


BUILDERS = [
    ("add8", lambda b: cm.adder_circuit(8, b), 2, lambda v: (v[0] + v[1]) % 256, 8),
    ("add16", lambda b: cm.adder_circuit(16, b), 2, lambda v: (v[0] + v[1]) % 65536, 16),
    ("lt32", lambda b: cm.compare_circuit(32, b), 2, lambda v: (v[0] < v[1]).astype(np.uint64), 32),
    ("mul8", lambda b: cm.multiply_circuit(8, b), 2, lambda v: (v[0] * v[1]) % 256, 8),
    ("mul16", lambda b: cm.multiply_circuit(16, b), 2, lambda v: (v[0] * v[1]) % 65536, 16),
    ("radd8", lambda b: cm.reduce_add_circuit(8, b), 2, lambda v: (v[0] + v[1]) % 256, 8),
    ("rmax8", lambda b: cm.reduce_max_circuit(8, b), 2, lambda v: np.maximum(v[0], v[1]), 8),
    ("pipe8", lambda b: cm.pipeline_circuit(8, b), 3, lambda v: np.where((v[0] + v[1]) % 256 < v[2], (v[0] + v[1]) % 256, v[2]), 8),
]


@pytest.mark.parametrize("basis", ["wire", "model", "fast"])
@pytest.mark.parametrize("name,build,nin,plain,bits", BUILDERS)
def test_circuit_builders_plain_semantics(name, build, nin, plain, bits, basis):
    c = build(basis)
    rng = np.random.default_rng(1)
    vals = [rng.integers(0, 2**bits, 200, dtype=np.uint64) for _ in range(nin)]
    bits_in = np.concatenate([cm.pack_words(v, bits) for v in vals], axis=1)
    out = po.Oracle().circuit_plain(c.to_c(), bits_in, c.out_bits)
    assert np.array_equal(cm.unpack_words(out), plain(vals).astype(np.uint64))
    if basis == "wire":                                   # only AND/OR/NOT may cross the herd wire (node.proto:25-30)
        assert all(op in (cm.Operation.AND, cm.Operation.OR, cm.Operation.NOT) for k, op, *_ in c.nodes if k == cm.NodeKind.OPERATION)


def test_select_circuit_plain():
    for basis in ("wire", "model"):
        c = cm.select_circuit(8, basis)
        rng = np.random.default_rng(2)
        s = rng.integers(0, 2, 100); a = rng.integers(0, 256, 100); b = rng.integers(0, 256, 100)
        bits_in = np.concatenate([s.reshape(-1, 1).astype(np.uint8), cm.pack_words(a, 8), cm.pack_words(b, 8)], axis=1)
        out = po.Oracle().circuit_plain(c.to_c(), bits_in, 8)
        assert np.array_equal(cm.unpack_words(out), np.where(s == 1, a, b))


def test_analyze_counts_and_level_structure():
    info = api.analyze(cm.adder_circuit(8))
    assert info["in_bits"] == 16 and info["out_bits"] == 8
    assert info["boot_gates"] == 64 and info["levels"] == 16            # 1 xor + 7 full adders (last without carry)
    m = api.analyze(cm.adder_circuit(8, "model"))
    assert (m["boot_gates"], m["levels"]) == (info["boot_gates"], info["levels"])   # XOR lowers to the same 3 bootstraps
    one = api.analyze(cm.single_gate_circuit())
    assert (one["boot_gates"], one["levels"], one["slots"]) == (1, 1, 1)
    big = api.analyze(cm.multiply_circuit(32))
    assert big["slots"] < big["boot_gates"] / 4                          # liveness reuse keeps the slot file small
    assert api.analyze(cm.compare_circuit(32))["boot_gates"] == 187


def test_fast_xor_nodes_cost_one_bootstrap_and_decrypt(test_key):
    """HERDB_NODE_FAST (include/herdb.h): an XOR node becomes one XOR_FAST bootstrap; the oracle's interpreter honours
    the same flag, and the encrypted result still decrypts to the plaintext sum"""
    wire, fast = api.analyze(cm.adder_circuit(8, "wire")), api.analyze(cm.adder_circuit(8, "fast"))
    assert wire["boot_gates"] == 64 and fast["boot_gates"] == 34 and fast["levels"] < wire["levels"]
    c = cm.adder_circuit(8, "fast")
    x = np.array([200, 13, 255]); y = np.array([100, 29, 255])
    bits = np.concatenate([cm.pack_words(x, 8), cm.pack_words(y, 8)], axis=1)
    rows = cs.encrypt_bits(test_key["n"], test_key["sk"], bits, seed=5)
    out = po.Oracle(test_key["blob"]).circuit_rows(c.to_c(), rows, 8)
    assert np.array_equal(cm.unpack_words(cs.decrypt_bits(test_key["n"], test_key["sk"], out)), (x + y) % 256)


def test_not_and_constants_cost_no_bootstrap():
    c = cm.Circuit([[cm.DataType.BIT]], [("o", cm.DataType.UINT8)])
    a = c.input(0, 0, 0)
    for b in range(8):
        c.output_bit(0, b, c.not_(c.not_(c.not_(a))) if b % 2 else c.constant(b % 3 == 0))
    info = api.analyze(c)
    assert info["boot_gates"] == 0 and info["levels"] == 0 and info["slots"] == 0


def _bad_circuits():
    D = cm.DataType
    c1 = cm.Circuit([[D.BIT, D.BIT]], [("o", D.BIT)])          # AND with one operand
    c1.output_bit(0, 0, c1.op(cm.Operation.AND, c1.input(0, 0, 0)))
    c2 = cm.Circuit([[D.BIT]], [("o", D.BIT)])                 # input field out of range
    c2.output_bit(0, 0, c2.input(0, 3, 0))
    c3 = cm.Circuit([[D.BIT]], [("o", D.UINT8)])               # output bit 1..7 never driven
    c3.output_bit(0, 0, c3.input(0, 0, 0))
    c4 = cm.Circuit([[D.BIT, D.BIT]], [("o", D.BIT)])          # cycle
    a = c4.input(0, 0, 0); g1 = c4.and_(a, a); g2 = c4.or_(g1, a); c4.edges.append((g2, g1)); c4.output_bit(0, 0, g2)
    c5 = cm.Circuit([[D.BIT]], [("o", D.BIT)])                 # bit index beyond the field width
    c5.output_bit(0, 0, c5.input(0, 0, 1))
    c6 = cm.Circuit([[D.BIT]], [("o", D.BIT)])                 # edge to a node that does not exist
    c6.output_bit(0, 0, c6.input(0, 0, 0)); c6.edges.append((0, 77))
    return [c1, c2, c3, c4, c5, c6]


@pytest.mark.parametrize("idx", range(6))
def test_malformed_circuits_are_rejected(idx):
    c = _bad_circuits()[idx]
    with pytest.raises(api.HerdbError) as e:
        api.analyze(c)
    assert e.value.code == -1
    if idx != 2:      # (the oracle does not check that every output bit is driven)
        with pytest.raises(ValueError):
            po.Oracle().circuit_plain(c.to_c(), np.zeros((1, max(1, c.in_bits)), np.uint8), max(1, c.out_bits))


def test_partition_file_framing(tmp_path):
    n = 32
    rows = np.random.default_rng(0).integers(0, 1024, (5, 16, n + 1)).astype(np.uint16)
    p = tmp_path / "0"
    api.partition_write(p, rows)
    raw = p.read_bytes()
    row_bytes = 16 * (n + 1) * 2
    assert len(raw) == 5 * (4 + row_bytes)
    assert int.from_bytes(raw[:4], "little") == row_bytes                # [u32 len][payload] (storage_service.cpp:19-28)
    assert raw[4:4 + row_bytes] == rows[0].tobytes()
    assert not (tmp_path / "0.tmp").exists()
    assert np.array_equal(api.partition_read(p, 5, 16, n), rows)
    for bad_rows, bad_bits in [(6, 16), (4, 16), (5, 15)]:               # too few rows, too many rows, ragged length
        with pytest.raises(api.HerdbError):
            api.partition_read(p, bad_rows, bad_bits, n)
    with pytest.raises(api.HerdbError):
        api.partition_read(tmp_path / "missing", 1, 16, n)
    api.partition_write(tmp_path / "empty", np.zeros((0, 16, n + 1), np.uint16))     # empty partition
    assert (tmp_path / "empty").read_bytes() == b""
    assert api.partition_read(tmp_path / "empty", 0, 16, n).shape[0] == 0


def test_partition_row_counts_match_herdsman():
    # storage_service.cpp:121-129: first rows % P partitions get one extra row
    for rows, parts in [(10, 3), (1024, 8), (7, 7), (3, 8), (0, 4), (1_000_000, 8)]:
        sizes = [api.partition_rows(rows, parts, k) for k in range(parts)]
        assert sum(sizes) == rows and max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
        from herdsman_b200.sharding import row_slice
        assert [row_slice(rows, parts, k)[1] - row_slice(rows, parts, k)[0] for k in range(parts)] == sizes


def test_task_handle_callback_semantics():
    """IWorkerGroup::TaskHandle: callback installed after scheduling still observes the completion"""
    h = api.TaskHandle()
    seen = []
    assert h.status() == api.TaskHandle.Status.PENDING and not h.completed()
    h._mark_completed(api.TaskHandle.Status.COMPLETED)
    h.set_completion_callback(seen.append)
    assert seen == [api.TaskHandle.Status.COMPLETED] and h.completed()
    h2 = api.TaskHandle(); h2.set_completion_callback(seen.append); h2._mark_completed(api.TaskHandle.Status.ERROR)
    assert seen[-1] == api.TaskHandle.Status.ERROR


def test_reference_side_adapter_compiles_against_herdsman_headers(tmp_path):
    """include/herd_b200/cuda_worker_group.hpp is the binding a herdsman maintainer adds; where the
    reference tree is present (build container only) it must compile against the real IWorkerGroup /
    MapTask / Circuit headers (libuuid's header is absent here: a 2-line typedef shim stands in)."""
    import subprocess
    ref = "/root/reference"
    if not os.path.isdir(os.path.join(ref, "include", "execution")):
        pytest.skip("reference tree not present on this machine")
    (tmp_path / "uuid").mkdir()
    (tmp_path / "uuid" / "uuid.h").write_text("#pragma once\ntypedef unsigned char uuid_t[16];\n")
    src = tmp_path / "check.cpp"
    src.write_text('#include "herd_b200/cuda_worker_group.hpp"\nint main() { return sizeof(CudaWorkerGroup) > 0 ? 0 : 1; }\n')
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    r = subprocess.run([cxx, "-std=c++20", "-fsyntax-only", "-Wall", "-I", str(tmp_path), "-I", ref + "/include",
                        "-I", ref + "/common/include", "-I", os.path.join(ROOT, "include"), str(src)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_partition_count_reads_to_the_end_of_the_file(tmp_path):
    """reduce inputs: InputDataFramePtr.row_count is only a hint (the reference labels hidden frames with
    row_count = partitions, execution_service.cpp:613-618), so a partition is read to its end"""
    rows = np.arange(5 * 3 * 33, dtype=np.uint16).reshape(5, 3, 33)
    api.partition_write(tmp_path / "p", rows)
    assert api.partition_count(tmp_path / "p", 3, 32) == 5
    api.partition_write(tmp_path / "e", rows[:0])
    assert api.partition_count(tmp_path / "e", 3, 32) == 0
    with pytest.raises(api.HerdbError):                          # other payload size
        api.partition_count(tmp_path / "p", 2, 32)
    raw = (tmp_path / "p").read_bytes()
    (tmp_path / "t").write_bytes(raw[:-7])                        # truncated last row
    with pytest.raises(api.HerdbError):
        api.partition_count(tmp_path / "t", 3, 32)
    with pytest.raises(api.HerdbError):
        api.partition_count(tmp_path / "missing", 3, 32)


def test_reference_reduce_tree_shapes():
    """tests/reduce_trees.py restates ReduceStageProgress (execution_service.cpp:560-689)"""
    import reduce_trees as rt
    seq = rt.build("SEQUENCED", 3, [4, 4, 3], "in", "out", "hid")
    assert seq == [([("in", 4, 0), ("in", 4, 1), ("in", 3, 2)], ("out", 1, 1))]
    par = rt.build("PARALLEL", 3, [4, 4, 3], "in", "out", "hid")
    assert len(par) == 4 and par[0] == ([("in", 4, 0)], ("hid", 3, 0))
    assert par[-1] == ([("hid", 3, 0), ("hid", 3, 1), ("hid", 3, 2)], ("out", 1, 1))      # label 3, one row each
    full = rt.build("PARALLEL_FULL", 5, [2] * 5, "in", "out", "hid", per_node_count=2)
    # 5 leaves -> 2 pairs + 1 carried = 3 -> 1 pair + 1 carried = 2 -> output
    assert len(full) == 5 + 2 + 1 + 1 and rt.hidden_partitions("PARALLEL_FULL", 5, 2) == 5 + 3 + 2
    assert full[5] == ([("hid", 1, 0), ("hid", 1, 1)], ("hid", 1, 5)) and full[7][0] == [("hid", 1, 5), ("hid", 1, 6)]
    assert full[-1] == ([("hid", 1, 7), ("hid", 1, 4)], ("out", 1, 1))
    produced = set()
    for ins, out in full:                                         # every input exists when its task runs
        assert all(n[0] == "in" or n in produced for n in ins)
        produced.add(out)
