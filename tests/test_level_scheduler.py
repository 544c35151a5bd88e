"""Level scheduler of the circuit compiler (herdsman_b200/csrc/circuit.cpp): gates with slack are moved to the
level whose last grid round they fill.  The schedule must stay a valid evaluation order with race-free slot
reuse (checked by running the compiled levels on plaintext bits, herdb_circuit_eval_plain), may never cost
more grid rounds than the as-soon-as-possible levels, and must not change the critical path."""
import numpy as np
import pytest

from herdsman_b200 import api, circuit as cm

WAVE = 148 * 4          # B200: 148 SMs x 4 ciphertexts per CTA

CIRCUITS = [
    ("add8", lambda: cm.adder_circuit(8), 2, 8, lambda v: (v[0] + v[1]) % 256),
    ("add8_fast", lambda: cm.adder_circuit(8, "fast"), 2, 8, lambda v: (v[0] + v[1]) % 256),
    ("lt32", lambda: cm.compare_circuit(32), 2, 32, lambda v: (v[0] < v[1]).astype(np.uint64)),
    ("mul8", lambda: cm.multiply_circuit(8), 2, 8, lambda v: (v[0] * v[1]) % 256),
    ("mul16", lambda: cm.multiply_circuit(16, "model"), 2, 16, lambda v: (v[0] * v[1]) % 65536),
    ("rmax8", lambda: cm.reduce_max_circuit(8), 2, 8, lambda v: np.maximum(v[0], v[1])),
    ("pipe8", lambda: cm.pipeline_circuit(8), 3, 8, lambda v: np.where((v[0] + v[1]) % 256 < v[2], (v[0] + v[1]) % 256, v[2])),
]
HINTS = [(0, 0), (1, WAVE), (37, WAVE), (148, WAVE), (256, WAVE), (1024, WAVE), (4096, WAVE), (100, 64), (5, 3)]


@pytest.mark.parametrize("name,build,nin,bits,plain", CIRCUITS)
def test_every_schedule_evaluates_the_circuit(name, build, nin, bits, plain):
    c = build()
    rng = np.random.default_rng(5)
    vals = [rng.integers(0, 2**bits, 12, dtype=np.uint64) for _ in range(nin)]
    bits_in = np.concatenate([cm.pack_words(v, bits) for v in vals], axis=1)
    want = plain(vals).astype(np.uint64)
    for rows, wave in HINTS:
        got = np.stack([api.eval_plain(c, bits_in[r], rows, wave) for r in range(bits_in.shape[0])])
        assert np.array_equal(cm.unpack_words(got).reshape(-1), want), (name, rows, wave)


@pytest.mark.parametrize("name,build,nin,bits,plain", CIRCUITS)
def test_schedule_never_costs_more_and_keeps_the_critical_path(name, build, nin, bits, plain):
    c = build()
    base = api.schedule(c)
    info = api.analyze(c)
    assert base["levels"] == info["levels"] and base["policy"] == 0 and sum(base["level_widths"]) == info["boot_gates"]
    for rows, wave in HINTS[1:]:
        s = api.schedule(c, rows, wave)
        assert sum(s["level_widths"]) == info["boot_gates"]
        assert s["levels"] <= info["levels"]                                   # never a longer dependency chain
        assert s["rounds"] <= s["rounds_asap"]
        rounds = sum(-(-w * rows // wave) for w in s["level_widths"])
        assert rounds == s["rounds"]
        assert sum(-(-w * rows // wave) for w in base["level_widths"]) == s["rounds_asap"]
        assert s["rounds"] >= -(-info["boot_gates"] * rows // wave)             # the bound no schedule beats


def test_known_gains():
    add8 = cm.adder_circuit(8)
    s = api.schedule(add8, 148, WAVE)               # 64 gates x 148 rows = 16 full rounds: every round filled
    assert (s["rounds_asap"], s["rounds"]) == (22, 16)
    s = api.schedule(add8, 1024, WAVE)              # the reference's 1,024-row case: 120 -> 114 rounds (bound 111)
    assert s["rounds_asap"] == 120 and s["rounds"] <= 114
    s = api.schedule(cm.multiply_circuit(32), 512, WAVE)
    assert s["rounds"] <= 4130 < s["rounds_asap"]
    one = api.schedule(cm.single_gate_circuit(), 33152, WAVE)
    assert (one["levels"], one["rounds"], one["policy"]) == (1, 56, 0)
