"""GPU tests of every column width herdsman defines (common/src/native_type_mapping.cpp:8-35: BIT 1,
(U)INT8 8, (U)INT16 16, (U)INT32 32, (U)INT64 64 bit-ciphertexts per field): one circuit per DataType
through the C ABI, bit-exact against the oracle's per-row interpreter at the n = 32 test set, plus a
mixed-width row so that field offsets inside a row are exercised."""
import numpy as np
import pytest

from herdsman_b200 import api, circuit as cm
import clientsim as cs

pytestmark = pytest.mark.gpu
D = cm.DataType


@pytest.fixture(scope="module")
def env(test_key, oracle_lib):
    b = api.Backend(devices=(0,))
    b.load_key_mem("w", test_key["blob"])
    yield {"be": b, "oracle": oracle_lib.Oracle(test_key["blob"]), **test_key}
    b.close()


def _bitwise(t):
    """(t x, t y) -> t: out_i = (x_i AND y_i) for even i, (x_i OR NOT y_i) for odd i"""
    c = cm.Circuit([[t, t]], [("o", t)])
    xs, ys = c.input_word(0, 0), c.input_word(0, 1)
    c.output_word(0, [c.and_(a, b) if i % 2 == 0 else c.or_(a, c.not_(b)) for i, (a, b) in enumerate(zip(xs, ys))])
    return c


def _plain_bitwise(x, y, bits):
    even = sum(1 << i for i in range(0, bits, 2)); odd = ((1 << bits) - 1) ^ even
    return ((x & y) & even) | ((x | (~y & ((1 << bits) - 1))) & odd)


@pytest.mark.parametrize("t", [D.BIT, D.UINT8, D.UINT16, D.UINT32, D.UINT64, D.INT8, D.INT16, D.INT32, D.INT64])
def test_every_data_type_width(env, t):
    e = env
    bits = cm.BIT_WIDTH[t]
    c = _bitwise(t)
    assert c.in_bits == 2 * bits and c.out_bits == bits
    rows = 5
    rng = np.random.default_rng(int(t))
    x = [int(v) for v in rng.integers(0, 1 << min(bits, 63), rows, dtype=np.uint64)]
    y = [int(v) for v in rng.integers(0, 1 << min(bits, 63), rows, dtype=np.uint64)]
    if bits == 64:
        x[0] |= 1 << 63; y[1] |= 1 << 63
    mat = np.array([[(v >> i) & 1 for i in range(bits)] + [(w >> i) & 1 for i in range(bits)] for v, w in zip(x, y)], np.uint8)
    rows_in = cs.encrypt_bits(e["n"], e["sk"], mat, seed=60 + int(t))
    out = e["be"].eval_circuit_rows("w", c, rows_in)
    assert out.shape == (rows, bits, e["n"] + 1)
    assert np.array_equal(out, e["oracle"].circuit_rows(c.to_c(), rows_in, bits))
    dec = cs.decrypt_bits(e["n"], e["sk"], out)
    got = [sum(int(dec[r, i]) << i for i in range(bits)) for r in range(rows)]
    assert got == [_plain_bitwise(v, w, bits) for v, w in zip(x, y)]


def test_wide_adders_16_and_64(env):
    """ripple-carry adders on UINT16 and UINT64 columns: 16 / 64 dependent levels, carries across the whole word"""
    e = env
    for bits, x, y in [(16, [0xFFFF, 0x1234, 0x8000], [1, 0xEDCC, 0x8000]),
                       (64, [0xFFFFFFFFFFFFFFFF, 0x0123456789ABCDEF, 1 << 63], [1, 0xFEDCBA9876543211, 1 << 63])]:
        c = cm.adder_circuit(bits)
        mat = np.array([[(v >> i) & 1 for i in range(bits)] + [(w >> i) & 1 for i in range(bits)] for v, w in zip(x, y)], np.uint8)
        rows_in = cs.encrypt_bits(e["n"], e["sk"], mat, seed=70 + bits)
        out = e["be"].eval_circuit_rows("w", c, rows_in)
        assert np.array_equal(out, e["oracle"].circuit_rows(c.to_c(), rows_in, bits))
        dec = cs.decrypt_bits(e["n"], e["sk"], out)
        assert [sum(int(dec[r, i]) << i for i in range(bits)) for r in range(len(x))] == [(v + w) % (1 << bits) for v, w in zip(x, y)]


def test_mixed_width_row_field_offsets(env):
    """a row of (BIT, INT16, UINT64, INT8) columns -> (UINT8, BIT, INT64): field offsets 0, 1, 17, 81 in, 0, 8, 9 out"""
    e = env
    c = cm.Circuit([[D.BIT, D.INT16, D.UINT64, D.INT8]], [("a", D.UINT8), ("b", D.BIT), ("c", D.INT64)])
    s = c.input(0, 0, 0)
    i16, u64, i8 = c.input_word(0, 1), c.input_word(0, 2), c.input_word(0, 3)
    c.output_word(0, [c.and_(s, b) for b in i8])                              # sel ? i8 : 0
    c.output_bit(1, 0, c.or_(i16[15], u64[63]))                               # sign bits
    c.output_word(2, [c.or_(u64[i], i16[i % 16]) if i < 32 else c.not_(u64[i]) for i in range(64)])
    assert c.in_bits == 1 + 16 + 64 + 8 and c.out_bits == 8 + 1 + 64
    rng = np.random.default_rng(77)
    rows = 4
    mat = rng.integers(0, 2, (rows, c.in_bits)).astype(np.uint8)
    rows_in = cs.encrypt_bits(e["n"], e["sk"], mat, seed=78)
    out = e["be"].eval_circuit_rows("w", c, rows_in)
    assert np.array_equal(out, e["oracle"].circuit_rows(c.to_c(), rows_in, c.out_bits))
    dec = cs.decrypt_bits(e["n"], e["sk"], out)
    sel, w16, w64, w8 = mat[:, 0:1], mat[:, 1:17], mat[:, 17:81], mat[:, 81:89]
    want = np.concatenate([sel & w8, (w16[:, 15:16] | w64[:, 63:64]),
                           np.concatenate([w64[:, :32] | np.tile(w16, (1, 2)), 1 - w64[:, 32:]], axis=1)], axis=1)
    assert np.array_equal(dec, want)
