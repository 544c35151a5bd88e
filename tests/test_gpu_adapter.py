"""GPU test of the reference-side binding: include/herd_b200/cuda_worker_group.hpp compiled and LINKED against the
reference's own IWorkerGroup::TaskHandle, task / circuit / DAG model and UUID (tests/adapter/fake_executor.cpp, built in
the build container from the sources under /root/reference; the binary travels to the GPU box), driven the way
Executor::schedule_task_on_worker does -- callback installed after schedule_task returns (executor.cpp:82-94).
Checks: every completion is delivered exactly once (including tasks that fail within microseconds and a task refused
inside schedule_task: the lost-completion window of SURVEY.md 3.3), and the output partitions equal the oracle's."""
import os
import subprocess
import uuid

import numpy as np
import pytest

from herdsman_b200 import api, circuit as cm
import clientsim as cs

pytestmark = pytest.mark.gpu
BIN = os.path.join(os.path.dirname(__file__), "adapter", "_ref", "fake_executor")


@pytest.mark.skipif(not os.path.exists(BIN), reason="tests/adapter/_ref/fake_executor is built where the reference tree is present (build())")
def test_adapter_linked_against_reference_delivers_every_completion(tmp_path, test_key, oracle_lib):
    k = test_key
    key_dir, storage = tmp_path / "keys", tmp_path / "storage"
    session, in_df, out_df = (str(uuid.uuid4()) for _ in range(3))
    (key_dir / session).mkdir(parents=True)
    (key_dir / session / "1.key").write_bytes(k["blob"].tobytes())
    (storage / session / in_df).mkdir(parents=True); (storage / session / out_df).mkdir(parents=True)
    n_rows, parts = 14, 4
    rng = np.random.default_rng(2)
    x = rng.integers(0, 256, n_rows); y = rng.integers(0, 256, n_rows)
    rows = cs.encrypt_bits(k["n"], k["sk"], np.concatenate([cm.pack_words(x, 8), cm.pack_words(y, 8)], axis=1), seed=12)
    start = 0
    for p in range(parts):
        cnt = api.partition_rows(n_rows, parts, p)
        api.partition_write(storage / session / in_df / str(p), rows[start:start + cnt]); start += cnt
    for rep in range(3):                                   # the race is a timing matter: a few runs
        r = subprocess.run([BIN, str(storage), str(key_dir), session, in_df, out_df, str(parts), str(n_rows), "1"],
                           capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout + r.stderr
        lines = dict(l.split() for l in r.stdout.strip().splitlines())
        assert lines == {**{"map%d" % p: "COMPLETED" for p in range(parts)}, "missing_partition": "ERROR", "malformed_circuit": "ERROR", "no_key": "ERROR"}
    got = np.concatenate([api.partition_read(storage / session / out_df / str(p), api.partition_rows(n_rows, parts, p), 8, k["n"]) for p in range(parts)])
    assert np.array_equal(cm.unpack_words(cs.decrypt_bits(k["n"], k["sk"], got)), (x + y) % 256)
    # the harness builds its adder with the reference's own Circuit / DAG API; same gates as adder_circuit(8) => same ciphertexts
    assert np.array_equal(got, oracle_lib.Oracle(k["blob"]).circuit_rows(cm.adder_circuit(8).to_c(), rows, 8))
