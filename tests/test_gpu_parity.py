"""GPU parity tests: the CUDA path (through the C ABI of include/herdb.h) against the CPU oracle on
the same seeded keys and ciphertexts.  Bar: BIT-EXACT ciphertext equality (integer work)."""
import numpy as np
import pytest

from herdsman_b200 import api, circuit as cm
import clientsim as cs

pytestmark = pytest.mark.gpu

GATES = ["OR", "AND", "NOR", "NAND", "XOR_FAST", "XNOR_FAST"]
PLAIN = {"OR": lambda a, b: a | b, "AND": lambda a, b: a & b, "NOR": lambda a, b: 1 - (a | b),
         "NAND": lambda a, b: 1 - (a & b), "XOR_FAST": lambda a, b: a ^ b, "XNOR_FAST": lambda a, b: 1 - (a ^ b)}


@pytest.fixture(scope="module")
def backend():
    b = api.Backend(devices=(0,))
    yield b
    b.close()


@pytest.fixture(scope="module")
def test_session(backend, test_key, oracle_lib):
    backend.load_key_mem("test", test_key["blob"])
    return {"name": "test", "oracle": oracle_lib.Oracle(test_key["blob"]), **test_key}


@pytest.fixture(scope="module")
def std_session(backend, std128_key, oracle_lib):
    backend.load_key_mem("std128", std128_key["blob"])
    return {"name": "std128", "oracle": oracle_lib.Oracle(std128_key["blob"]), **std128_key}


def _inputs(sess, batch, seed):
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 2, batch).astype(np.uint8); b = rng.integers(0, 2, batch).astype(np.uint8)
    ca = cs.encrypt_bits(sess["n"], sess["sk"], a, seed=seed, first=0)
    cb = cs.encrypt_bits(sess["n"], sess["sk"], b, seed=seed, first=1 << 20)
    return a, b, ca, cb


@pytest.mark.parametrize("gate", GATES)
def test_gate_bit_exact_test_params(backend, test_session, gate):
    s = test_session
    a, b, ca, cb = _inputs(s, 37, seed=11)          # 37: ragged last group of 4
    out = backend.eval_gate_batch(s["name"], gate, ca, cb)
    ref = s["oracle"].gate(gate, ca, cb)
    assert np.array_equal(out, ref)
    assert np.array_equal(cs.decrypt_bits(s["n"], s["sk"], out), PLAIN[gate](a, b))


@pytest.mark.parametrize("batch", [1, 2, 3, 4, 5, 592, 593, 1200])
def test_gate_batch_sizes(backend, test_session, batch):
    s = test_session
    a, b, ca, cb = _inputs(s, batch, seed=batch)
    out = backend.eval_gate_batch(s["name"], "NAND", ca, cb)
    assert np.array_equal(cs.decrypt_bits(s["n"], s["sk"], out), 1 - (a & b))
    k = min(batch, 24)
    assert np.array_equal(out[:k], s["oracle"].gate("NAND", ca[:k], cb[:k]))
    assert np.array_equal(out[-k:], s["oracle"].gate("NAND", ca[-k:], cb[-k:]))


@pytest.mark.parametrize("batch", [2368, 2369, 3001])
def test_gate_batch_pipelined_pieces(backend, test_session, batch):
    """from four grid rounds up the host-buffer entry point cuts the batch into head / middle / tail pieces whose
    copies ride a second stream (backend.cpp: herdb_eval_gate_batch); every ciphertext must still equal the oracle's"""
    s = test_session
    a, b, ca, cb = _inputs(s, batch, seed=batch)
    out = backend.eval_gate_batch(s["name"], "OR", ca, cb)
    assert np.array_equal(cs.decrypt_bits(s["n"], s["sk"], out), a | b)
    assert np.array_equal(out, s["oracle"].gate("OR", ca, cb))


def test_gate_empty_batch(backend, test_session):
    s = test_session
    z = np.zeros((0, s["n"] + 1), np.uint16)
    assert backend.eval_gate_batch(s["name"], "AND", z, z).shape == (0, s["n"] + 1)


@pytest.mark.parametrize("gate", GATES)
def test_gate_bit_exact_std128(backend, std_session, gate):
    """every gate code of the C ABI at the production parameter set (n = 512), bit-exact with the oracle"""
    s = std_session
    a, b, ca, cb = _inputs(s, 24, seed=5)
    out = backend.eval_gate_batch(s["name"], gate, ca, cb)
    ref = s["oracle"].gate(gate, ca, cb)
    assert np.array_equal(out, ref), gate
    assert np.array_equal(cs.decrypt_bits(s["n"], s["sk"], out), PLAIN[gate](a, b))


def test_std128_large_batch_decrypts(backend, std_session):
    """size-independent property at a larger size: every output decrypts to NAND of the inputs and
    a bootstrapped output can feed another gate (noise is refreshed)"""
    s = std_session
    a, b, ca, cb = _inputs(s, 2048, seed=9)
    out = backend.eval_gate_batch(s["name"], "NAND", ca, cb)
    d1 = cs.decrypt_bits(s["n"], s["sk"], out)
    assert np.array_equal(d1, 1 - (a & b))
    out2 = backend.eval_gate_batch(s["name"], "OR", out, ca)
    assert np.array_equal(cs.decrypt_bits(s["n"], s["sk"], out2), d1 | a)


@pytest.mark.parametrize("name,build,bits_in,plain", [
    ("add8", lambda: cm.adder_circuit(8), (8, 8), lambda x, y: (x + y) % 256),
    ("add8_model", lambda: cm.adder_circuit(8, "model"), (8, 8), lambda x, y: (x + y) % 256),
    ("add8_fast", lambda: cm.adder_circuit(8, "fast"), (8, 8), lambda x, y: (x + y) % 256),      # XOR nodes as one XOR_FAST bootstrap
    ("lt8", lambda: cm.compare_circuit(8), (8, 8), lambda x, y: (x < y).astype(np.uint64)),
    ("mul8", lambda: cm.multiply_circuit(8), (8, 8), lambda x, y: (x * y) % 256),
    ("mul8_fast", lambda: cm.multiply_circuit(8, "fast"), (8, 8), lambda x, y: (x * y) % 256),
])
def test_circuit_bit_exact_test_params(backend, test_session, name, build, bits_in, plain):
    s = test_session
    c = build()
    rows = 9
    rng = np.random.default_rng(3)
    x = rng.integers(0, 256, rows); y = rng.integers(0, 256, rows)
    bits = np.concatenate([cm.pack_words(x, 8), cm.pack_words(y, 8)], axis=1)
    rows_in = cs.encrypt_bits(s["n"], s["sk"], bits, seed=21)
    out = backend.eval_circuit_rows(s["name"], c, rows_in)
    dec = cs.decrypt_bits(s["n"], s["sk"], out)
    assert np.array_equal(cm.unpack_words(dec), plain(x, y))
    ref = s["oracle"].circuit_rows(c.to_c(), rows_in, c.out_bits)
    assert np.array_equal(out, ref)


def test_circuit_constants_and_not_outputs(backend, test_session):
    """outputs fed directly by NOT / constants / inputs (no bootstrap on the way)"""
    s = test_session
    c = cm.Circuit([[cm.DataType.BIT, cm.DataType.BIT]], [("o", cm.DataType.UINT8)])
    a, b = c.input(0, 0, 0), c.input(0, 1, 0)
    one, zero = c.constant(1), c.constant(0)
    c.output_bit(0, 0, a); c.output_bit(0, 1, c.not_(a)); c.output_bit(0, 2, one); c.output_bit(0, 3, c.not_(one))
    c.output_bit(0, 4, c.and_(a, one)); c.output_bit(0, 5, c.or_(c.not_(b), zero)); c.output_bit(0, 6, c.nand(a, b))
    c.output_bit(0, 7, c.mux(a, b, c.not_(b)))
    bits = np.array([[0, 0], [0, 1], [1, 0], [1, 1]], np.uint8)
    rows_in = cs.encrypt_bits(s["n"], s["sk"], bits, seed=33)
    out = backend.eval_circuit_rows(s["name"], c, rows_in)
    ref = s["oracle"].circuit_rows(c.to_c(), rows_in, 8)
    assert np.array_equal(out, ref)
    dec = cs.decrypt_bits(s["n"], s["sk"], out)
    A, B = bits[:, 0], bits[:, 1]
    want = np.stack([A, 1 - A, np.ones(4, int), np.zeros(4, int), A, 1 - B, 1 - (A & B), np.where(A == 1, B, 1 - B)], axis=1)
    assert np.array_equal(dec, want)


def test_add8_std128_rows(backend, std_session):
    s = std_session
    c = cm.adder_circuit(8)
    rows = 64
    rng = np.random.default_rng(1)
    x = rng.integers(0, 256, rows); y = rng.integers(0, 256, rows)
    bits = np.concatenate([cm.pack_words(x, 8), cm.pack_words(y, 8)], axis=1)
    rows_in = cs.encrypt_bits(s["n"], s["sk"], bits, seed=1)
    out = backend.eval_circuit_rows(s["name"], c, rows_in)
    assert np.array_equal(cm.unpack_words(cs.decrypt_bits(s["n"], s["sk"], out)), (x + y) % 256)
    ref = s["oracle"].circuit_rows(c.to_c(), rows_in[:16], c.out_bits)     # oracle: ~3 s per row per core, OpenMP over rows
    assert np.array_equal(out[:16], ref)


@pytest.mark.parametrize("name,build,plain", [
    ("add8", lambda: cm.adder_circuit(8), lambda x, y: (x + y) % 256),
    ("mul8", lambda: cm.multiply_circuit(8), lambda x, y: (x * y) % 256),
])
def test_rescheduled_levels_bit_exact(backend, test_session, name, build, plain):
    """148 rows: the level scheduler (circuit.cpp) moves gates with slack off their as-soon-as-possible level
    to fill grid rounds; every gate is deterministic, so the ciphertexts must not change"""
    import torch
    s = test_session
    c = build()
    rows = 148
    wave = torch.cuda.get_device_properties(0).multi_processor_count * 4
    sched = api.schedule(c, rows, wave)
    assert sched["policy"] != 0 and sched["rounds"] < sched["rounds_asap"]       # this test is about a moved schedule
    rng = np.random.default_rng(8)
    x = rng.integers(0, 256, rows); y = rng.integers(0, 256, rows)
    bits = np.concatenate([cm.pack_words(x, 8), cm.pack_words(y, 8)], axis=1)
    rows_in = cs.encrypt_bits(s["n"], s["sk"], bits, seed=77)
    out = backend.eval_circuit_rows(s["name"], c, rows_in)
    assert np.array_equal(cm.unpack_words(cs.decrypt_bits(s["n"], s["sk"], out)), plain(x, y))
    ref = s["oracle"].circuit_rows(c.to_c(), rows_in[:6], c.out_bits)
    assert np.array_equal(out[:6], ref)
    small = backend.eval_circuit_rows(s["name"], c, rows_in[:5])                  # 5 rows: as-soon-as-possible levels
    assert np.array_equal(small, out[:5])


def test_small_memory_budget_chunks_rows_and_batches(backend, test_session, monkeypatch):
    """HERDB_MEM_BUDGET_MB bounds the scratch (slot file + extraction buffer): with 1 MiB an 8-bit adder runs 13 rows
    per chunk and a gate batch 511 gates per launch.  Same ciphertexts as the unchunked run."""
    s = test_session
    monkeypatch.setenv("HERDB_MEM_BUDGET_MB", "1")
    small = api.Backend(devices=(0,))
    try:
        small.load_key_mem("tiny", s["blob"])
        c = cm.adder_circuit(8)
        rows = 40
        rng = np.random.default_rng(12)
        x = rng.integers(0, 256, rows); y = rng.integers(0, 256, rows)
        bits = np.concatenate([cm.pack_words(x, 8), cm.pack_words(y, 8)], axis=1)
        rows_in = cs.encrypt_bits(s["n"], s["sk"], bits, seed=5)
        small.reset_stats()
        got = small.eval_circuit_rows("tiny", c, rows_in)
        assert small.stats()["blind_rotate_launches"] > api.analyze(c)["levels"]        # really chunked
        assert np.array_equal(got, backend.eval_circuit_rows(s["name"], c, rows_in))
        assert np.array_equal(cm.unpack_words(cs.decrypt_bits(s["n"], s["sk"], got)), (x + y) % 256)
        a, b, ca, cb = _inputs(s, 2500, seed=3)
        small.reset_stats()
        out = small.eval_gate_batch("tiny", "NAND", ca, cb)
        assert small.stats()["blind_rotate_launches"] >= 5
        assert np.array_equal(out, backend.eval_gate_batch(s["name"], "NAND", ca, cb))
        assert np.array_equal(cs.decrypt_bits(s["n"], s["sk"], out), 1 - (a & b))
    finally:
        small.close()


def test_concurrent_callers_share_a_device(backend, test_session):
    """the seam allows calls from several threads (executor thread + bench drivers): a device is driven by one host
    thread at a time (Device::mu), results are what each caller would get alone"""
    import threading
    s = test_session
    c = cm.adder_circuit(8)
    rng = np.random.default_rng(14)
    x = rng.integers(0, 256, 12); y = rng.integers(0, 256, 12)
    rows_in = cs.encrypt_bits(s["n"], s["sk"], np.concatenate([cm.pack_words(x, 8), cm.pack_words(y, 8)], axis=1), seed=6)
    a, b, ca, cb = _inputs(s, 700, seed=15)
    want_rows = backend.eval_circuit_rows(s["name"], c, rows_in)
    want_gate = backend.eval_gate_batch(s["name"], "AND", ca, cb)
    got, errs = {}, []

    def worker(k):
        try:
            for rep in range(3):
                if k % 2:
                    got[(k, rep)] = ("rows", backend.eval_circuit_rows(s["name"], c, rows_in))
                else:
                    got[(k, rep)] = ("gate", backend.eval_gate_batch(s["name"], "AND", ca, cb))
        except Exception as e:      # noqa: BLE001
            errs.append(e)

    th = [threading.Thread(target=worker, args=(k,)) for k in range(4)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errs and len(got) == 12
    for kind, v in got.values():
        assert np.array_equal(v, want_rows if kind == "rows" else want_gate)
