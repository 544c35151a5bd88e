"""GPU test of herdb_eval_gate_batch_dev_noalloc (include/herdb.h): the single-gate batch with its scratch in a fixed
array of the library's device image instead of cudaMalloc'ed buffers -- no counterpart in the reference, added
for callers that cannot allocate.

Through the C ABI, against the existing entry point herdb_eval_gate_batch_dev on the same device buffers: BIT-EXACT
(integer work: the tolerance of every parity test here is zero), on two calls in a row that share the scratch array
(the second one smaller and with other inputs, so stale contents of the first would show), plus the size limit.
Collected last on purpose (a failure here hides nothing after it)."""
import numpy as np
import pytest

from herdsman_b200 import api
import clientsim as cs

pytestmark = pytest.mark.gpu


def _inputs(key, batch, seed):
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 2, batch).astype(np.uint8); b = rng.integers(0, 2, batch).astype(np.uint8)
    ca = cs.encrypt_bits(key["n"], key["sk"], a, seed=seed, first=0)
    cb = cs.encrypt_bits(key["n"], key["sk"], b, seed=seed, first=1 << 20)
    return a, b, ca, cb


def test_no_alloc_entry_point_equals_the_allocating_one_on_two_calls_in_a_row(test_key, oracle_lib):
    import torch
    k = test_key
    be = api.Backend(devices=(0,))
    try:
        be.load_key_mem("na", k["blob"])
        # every device buffer the calls touch exists before the first call: the entry point itself allocates nothing
        calls = [("NAND", 1000, 101), ("XOR_FAST", 37, 202)]             # body + tail shapes; then a small ragged batch
        bufs = []
        for gate, batch, seed in calls:
            a, b, ca, cb = _inputs(k, batch, seed)
            d_a = torch.from_numpy(ca).cuda(); d_b = torch.from_numpy(cb).cuda()
            bufs.append((gate, batch, a, b, ca, cb, d_a, d_b, torch.zeros_like(d_a), torch.zeros_like(d_a)))
        torch.cuda.synchronize()
        launches0 = be.stats()["kernel_launches"]
        for gate, batch, a, b, ca, cb, d_a, d_b, d_new, d_old in bufs:     # the two no-alloc calls, back to back
            be.eval_gate_batch_dev_noalloc("na", gate, d_a, d_b, d_new, batch)
        assert be.stats()["kernel_launches"] > launches0                  # our kernels ran (no fallback of any kind)
        plain = {"NAND": lambda x, y: 1 - (x & y), "XOR_FAST": lambda x, y: x ^ y}
        orc = oracle_lib.Oracle(k["blob"])
        for gate, batch, a, b, ca, cb, d_a, d_b, d_new, d_old in bufs:
            be.eval_gate_batch_dev("na", gate, d_a, d_b, d_old, batch)    # the existing entry point, same inputs
            new, old = d_new.cpu().numpy(), d_old.cpu().numpy()
            assert np.array_equal(new, old), gate                         # bit-exact
            assert np.array_equal(cs.decrypt_bits(k["n"], k["sk"], new), plain[gate](a, b))
            m = min(batch, 12)
            assert np.array_equal(new[:m], orc.gate(gate, ca[:m], cb[:m]))
        # limits: the maximum itself is accepted, one more is refused with a clear message, nothing is written
        B = api.NOALLOC_MAX_BATCH
        a, b, ca, cb = _inputs(k, B + 1, 303)
        d_a = torch.from_numpy(ca).cuda(); d_b = torch.from_numpy(cb).cuda(); d_out = torch.zeros_like(d_a)
        with pytest.raises(api.HerdbError) as e:
            be.eval_gate_batch_dev_noalloc("na", "AND", d_a, d_b, d_out, B + 1)
        assert e.value.code == -1 and "HERDB_NOALLOC_MAX_BATCH" in str(e.value)
        assert int(d_out.count_nonzero()) == 0
        be.eval_gate_batch_dev_noalloc("na", "AND", d_a, d_b, d_out, B)
        assert np.array_equal(cs.decrypt_bits(k["n"], k["sk"], d_out[:B].cpu().numpy()), a[:B] & b[:B])
        be.eval_gate_batch_dev_noalloc("na", "AND", d_a, d_b, d_out, 0)   # empty batch: nothing to do
        with pytest.raises(api.HerdbError):
            be.eval_gate_batch_dev_noalloc("no-such-session", "AND", d_a, d_b, d_out, 4)
    finally:
        be.close()


def test_no_alloc_entry_point_std128(std128_key):
    """the production parameter set, and a fresh context (its own one-time address lookup) after the first test's"""
    import torch
    k = std128_key
    be = api.Backend(devices=(0,))
    try:
        be.load_key_mem("na512", k["blob"])
        for batch, seed in [(300, 11), (9, 12)]:
            a, b, ca, cb = _inputs(k, batch, seed)
            d_a = torch.from_numpy(ca).cuda(); d_b = torch.from_numpy(cb).cuda()
            d_new = torch.zeros_like(d_a); d_old = torch.zeros_like(d_a)
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info()[0]
            be.eval_gate_batch_dev_noalloc("na512", "OR", d_a, d_b, d_new, batch)
            if seed == 12:                       # second call (module image and scratch array are resident by now):
                assert torch.cuda.mem_get_info()[0] == free_before      # the call took no device memory
            be.eval_gate_batch_dev("na512", "OR", d_a, d_b, d_old, batch)
            assert np.array_equal(d_new.cpu().numpy(), d_old.cpu().numpy())
            assert np.array_equal(cs.decrypt_bits(k["n"], k["sk"], d_new.cpu().numpy()), a | b)
    finally:
        be.close()
