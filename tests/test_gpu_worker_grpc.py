"""GPU test of the gRPC `Worker` façade: a client that does what GrpcWorkerGroup does -- serialise
the task with the reference's wire schema, call /herd.proto.Worker/map and /reduce on an insecure channel
(grpc_worker_group.cpp:70-105) -- against herdsman_b200.worker_server backed by the CUDA backend.  The output
partitions must equal the oracle's, and failures must come back as non-OK statuses."""
import uuid

import grpc
import numpy as np
import pytest

from herdsman_b200 import api, circuit as cm, worker_proto as wp, worker_server as ws
import clientsim as cs

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def served(tmp_path_factory, test_key, oracle_lib):
    root = tmp_path_factory.mktemp("grpc")
    key_dir, storage = root / "keys", root / "storage"
    session = str(uuid.uuid4())
    (key_dir / session).mkdir(parents=True)
    (key_dir / session / "1.key").write_bytes(test_key["blob"].tobytes())
    wg = api.CudaWorkerGroup(str(storage), str(key_dir), devices=(0,))
    server, port = ws.serve(wg, "127.0.0.1:0")
    stub = ws.WorkerStub("127.0.0.1:%d" % port)
    yield {"stub": stub, "session": session, "storage": storage, "oracle": oracle_lib.Oracle(test_key["blob"]), **test_key}
    stub.close(); server.stop(0); wg.close()


def _df(w):
    df = str(uuid.uuid4()); (w["storage"] / w["session"] / df).mkdir(parents=True)
    return df


def test_map_over_grpc_writes_the_oracles_partition(served):
    w = served
    c = cm.adder_circuit(8)                                   # AND / OR / NOT only: what the wire can carry (node.proto:25-30)
    rng = np.random.default_rng(0)
    n_rows, parts = 9, 2
    x = rng.integers(0, 256, n_rows); y = rng.integers(0, 256, n_rows)
    rows = cs.encrypt_bits(w["n"], w["sk"], np.concatenate([cm.pack_words(x, 8), cm.pack_words(y, 8)], axis=1), seed=5)
    in_df, out_df = _df(w), _df(w)
    start = 0
    for k in range(parts):
        cnt = api.partition_rows(n_rows, parts, k)
        api.partition_write(w["storage"] / w["session"] / in_df / str(k), rows[start:start + cnt]); start += cnt
    futures = []
    for k in range(parts):                                    # Asyncmap: both partitions in flight
        t = api.MapTask(w["session"], api.InputDataFramePtr(api.DataFramePtr(in_df, k), api.partition_rows(n_rows, parts, k)),
                        api.DataFramePtr(out_df, k), api.CryptoKeyPtr(1), c)
        futures.append(w["stub"].map.future(wp.MapTask.FromString(ws.task_to_proto(t).SerializeToString()), timeout=300))
    for f in futures:
        assert f.result() == wp.Empty()                       # returns when the partition is written
    got = np.concatenate([api.partition_read(w["storage"] / w["session"] / out_df / str(k), api.partition_rows(n_rows, parts, k), 8, w["n"]) for k in range(parts)])
    assert np.array_equal(got, w["oracle"].circuit_rows(c.to_c(), rows, 8))
    assert np.array_equal(cm.unpack_words(cs.decrypt_bits(w["n"], w["sk"], got)), (x + y) % 256)


def test_reduce_over_grpc_and_error_statuses(served):
    w = served
    c = cm.reduce_add_circuit(8)
    vals = np.random.default_rng(1).integers(0, 256, 5)
    rows = cs.encrypt_bits(w["n"], w["sk"], cm.pack_words(vals, 8), seed=6)
    in_df, out_df = _df(w), _df(w)
    api.partition_write(w["storage"] / w["session"] / in_df / "0", rows[:3]); api.partition_write(w["storage"] / w["session"] / in_df / "1", rows[3:])
    t = api.ReduceTask(w["session"], [api.InputDataFramePtr(api.DataFramePtr(in_df, 0), 3), api.InputDataFramePtr(api.DataFramePtr(in_df, 1), 2)],
                       api.DataFramePtr(out_df, 1), api.CryptoKeyPtr(1), c)
    assert w["stub"].reduce(ws.task_to_proto(t), timeout=300) == wp.Empty()
    got = api.partition_read(w["storage"] / w["session"] / out_df / "1", 1, 8, w["n"])
    assert cm.unpack_words(cs.decrypt_bits(w["n"], w["sk"], got))[0] == vals.sum() % 256
    # missing partition => non-OK (the reference client turns it into Status::ERROR and fails the job)
    bad = api.MapTask(w["session"], api.InputDataFramePtr(api.DataFramePtr(in_df, 9), 1), api.DataFramePtr(out_df, 0), api.CryptoKeyPtr(1), cm.adder_circuit(8))
    with pytest.raises(grpc.RpcError) as e:
        w["stub"].map(ws.task_to_proto(bad), timeout=120)
    assert e.value.code() == grpc.StatusCode.INTERNAL and "partition" in e.value.details()
    nokey = api.MapTask(str(uuid.uuid4()), bad.input_data_frame_ptr, bad.output_data_frame_ptr, api.CryptoKeyPtr(1), cm.adder_circuit(8))
    with pytest.raises(grpc.RpcError):
        w["stub"].map(ws.task_to_proto(nokey), timeout=120)
