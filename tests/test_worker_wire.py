"""CPU tests of the gRPC `Worker` façade (herdsman_b200/worker_proto.py, worker_server.py): the hand-built
descriptors against the reference's .proto files, hand-encoded wire bytes, model <-> proto round trips, and the
RPC plumbing with a recording worker group (no GPU needed).  The GPU end of it is tests/test_gpu_worker_grpc.py."""
import os
import re
import threading

import numpy as np
import pytest

from herdsman_b200 import api, circuit as cm, worker_proto as wp, worker_server as ws

REF_PROTO = "/root/reference/proto/src"


def _parse_proto(path):
    """{message: {field: (number, type, repeated)}}, {enum: {name: number}}, {service: {rpc: (in, out)}} of one proto3 file"""
    text = re.sub(r"//[^\n]*", "", open(path).read())
    msgs, enums, svcs = {}, {}, {}
    for kind, name, body in re.findall(r"\b(message|enum|service)\s+(\w+)\s*\{((?:[^{}]|\{[^{}]*\})*)\}", text):
        if kind == "enum":
            enums[name] = {k: int(v) for k, v in re.findall(r"(\w+)\s*=\s*(\d+)\s*;", body)}
        elif kind == "service":
            svcs[name] = {r: (i, o) for r, i, o in re.findall(r"rpc\s+(\w+)\s*\(\s*(\w+)\s*\)\s*returns\s*\(\s*(\w+)\s*\)", body)}
        else:
            flat = re.sub(r"oneof\s+\w+\s*\{([^{}]*)\}", r"\1", body)
            msgs[name] = {f: (int(n), t, bool(rep)) for rep, t, f, n in re.findall(r"(repeated\s+)?([\w.]+)\s+(\w+)\s*=\s*(\d+)\s*;", flat)}
    return msgs, enums, svcs


@pytest.mark.skipif(not os.path.isdir(REF_PROTO), reason="reference tree not present on this machine")
def test_descriptors_mirror_the_reference_proto_files_field_for_field():
    from google.protobuf.descriptor import FieldDescriptor as FD
    scalar = {FD.TYPE_UINT32: "uint32", FD.TYPE_UINT64: "uint64", FD.TYPE_STRING: "string", FD.TYPE_BOOL: "bool", FD.TYPE_BYTES: "bytes"}
    for fd in wp.FILES:
        msgs, enums, svcs = _parse_proto(os.path.join(REF_PROTO, fd.name))
        ours_msgs = {m.name: m for m in fd.message_type}
        assert set(ours_msgs) == set(msgs), fd.name
        for name, fields in msgs.items():
            got = {}
            for f in ours_msgs[name].field:
                t = scalar.get(f.type) or f.type_name.rsplit(".", 1)[1]
                got[f.name] = (f.number, t, f.label == f.LABEL_REPEATED)
            assert got == fields, (fd.name, name)
        assert {e.name: {v.name: v.number for v in e.value} for e in fd.enum_type} == enums, fd.name
        assert {s.name: {m.name: (m.input_type.rsplit(".", 1)[1], m.output_type.rsplit(".", 1)[1]) for m in s.method} for s in fd.service} == svcs
        assert fd.package == "herd.proto" and fd.syntax == "proto3"
    node = next(m for fd in wp.FILES for m in fd.message_type if m.name == "Node")
    assert [f.oneof_index for f in node.field] == [0, 0, 0, 0] and node.oneof_decl[0].name == "node"


def _varint(v):
    out = bytearray()
    while True:
        b = v & 0x7F; v >>= 7
        out.append(b | (0x80 if v else 0))
        if not v:
            return bytes(out)


def _ld(field, payload):
    return _varint(field << 3 | 2) + _varint(len(payload)) + payload


def _vi(field, v):
    return _varint(field << 3 | 0) + _varint(v)


def test_hand_encoded_map_task_bytes_parse_into_the_model():
    """wire bytes written out by hand from the reference's field numbers (worker.proto:11-37, circuit.proto, node.proto,
    common.proto) -- independent of the descriptors -- for: out = a AND NOT b on two BIT columns"""
    sess, din, dout = "0a1b2c3d-0000-4000-8000-000000000001", "0a1b2c3d-0000-4000-8000-000000000002", "0a1b2c3d-0000-4000-8000-000000000003"
    nodes = [_ld(1, _vi(2, 0)),                         # Node.input{tuple 0 (default, omitted), field_index 0 omitted...}: field_index=0 -> empty; keep bit omitted
             _ld(1, _vi(2, 1)),                         # input field 1
             _ld(4, _vi(1, 2)),                         # operation NOT
             _ld(4, b""),                               # operation AND (enum 0: omitted on the wire)
             _ld(2, b"")]                               # output field 0 bit 0
    edges = [(0, 3), (1, 2), (2, 3), (3, 4)]
    circuit = _ld(1, _ld(1, bytes([0, 0]))) + _ld(2, _vi(1, 0) + _ld(2, b"o")) \
        + b"".join(_ld(3, n) for n in nodes) + b"".join(_ld(4, (_vi(1, s) if s else b"") + _vi(2, e)) for s, e in edges)
    msg = _ld(1, sess.encode()) + _ld(2, _ld(1, _ld(1, din.encode()) + _vi(2, 3)) + _vi(4, 1000)) \
        + _ld(3, _ld(1, dout.encode()) + _vi(2, 3)) + _ld(4, b"") + _ld(5, circuit)
    t = ws.map_task_to_model(wp.MapTask.FromString(msg))
    assert (t.session_uuid, t.input_data_frame_ptr.pointer.uuid, t.input_data_frame_ptr.pointer.partition, t.input_data_frame_ptr.row_count) == (sess, din, 3, 1000)
    assert (t.output_data_frame_ptr.uuid, t.output_data_frame_ptr.partition, t.crypto_key_ptr.schema_type) == (dout, 3, 1)   # BINFHE(0) -> model 1 -> 1.key
    c = t.circuit
    assert c.inputs == [[cm.DataType.BIT, cm.DataType.BIT]] and c.output == [("o", cm.DataType.BIT)]
    K, O = cm.NodeKind, cm.Operation
    assert [(n[0], n[1], n[4]) for n in c.nodes] == [(K.INPUT, 0, 0), (K.INPUT, 0, 1), (K.OPERATION, O.NOT, 0), (K.OPERATION, O.AND, 0), (K.OUTPUT, 0, 0)]
    assert c.edges == edges
    from oracle import pyoracle as po
    bits = np.array([[0, 0], [0, 1], [1, 0], [1, 1]], np.uint8)
    assert po.Oracle().circuit_plain(c.to_c(), bits, 1).ravel().tolist() == [0, 0, 1, 0]
    # and our serializer emits the canonical proto3 bytes for the same task
    assert ws.task_to_proto(t).SerializeToString(deterministic=True) == wp.MapTask.FromString(msg).SerializeToString(deterministic=True)


@pytest.mark.parametrize("build", [lambda: cm.adder_circuit(8), lambda: cm.compare_circuit(32), lambda: cm.reduce_add_circuit(8)])
def test_circuit_round_trip_keeps_semantics_and_operand_order(build):
    from oracle import pyoracle as po
    c = build()
    pb = wp.Circuit.FromString(ws.circuit_to_proto(c).SerializeToString())
    c2 = ws.circuit_to_model(pb)
    assert c2.inputs == c.inputs and c2.output == c.output and [n[:6] for n in c2.nodes] == [n[:6] for n in c.nodes]
    assert sorted(c2.edges) == sorted(c.edges)
    bits = np.random.default_rng(0).integers(0, 2, (50, c.in_bits)).astype(np.uint8)
    assert np.array_equal(po.Oracle().circuit_plain(c2.to_c(), bits, c.out_bits), po.Oracle().circuit_plain(c.to_c(), bits, c.out_bits))
    assert api.analyze(c2) == api.analyze(c)


def test_operations_the_wire_cannot_carry_are_refused():
    with pytest.raises(ws.MappingError):                 # to_proto(Operation) throws for XOR / NAND / MUX (executor.cpp:242-245)
        ws.circuit_to_proto(cm.adder_circuit(8, "model"))
    pb = ws.circuit_to_proto(cm.single_gate_circuit(cm.Operation.AND))
    pb.nodes[2].operation.type = 5
    with pytest.raises(ws.MappingError):
        ws.circuit_to_model(pb)
    pb = ws.circuit_to_proto(cm.single_gate_circuit(cm.Operation.AND))
    pb.edges[0].end = 99
    with pytest.raises(ws.MappingError):
        ws.circuit_to_model(pb)


class _RecordingGroup:
    """IWorkerGroup stand-in: completes tasks on another thread, ERROR for session 'bad'"""

    class backend:                                        # noqa: N801
        class L:                                          # noqa: N801
            @staticmethod
            def herdb_last_error(_):
                return b"recorded failure"
        h = None

    def __init__(self):
        self.tasks = []

    def concurrent_workers(self):
        return 2

    def schedule_task(self, task):
        self.tasks.append(task)
        h = api.TaskHandle()
        st = api.TaskHandle.Status.ERROR if task.session_uuid == "bad" else api.TaskHandle.Status.COMPLETED
        threading.Timer(0.01, h._mark_completed, args=(st,)).start()
        return h


def test_rpc_plumbing_map_and_reduce_over_a_channel():
    import grpc
    wg = _RecordingGroup()
    server, port = ws.serve(wg, "127.0.0.1:0")
    stub = ws.WorkerStub("127.0.0.1:%d" % port)
    try:
        c = cm.adder_circuit(8)
        mt = api.MapTask("s-1", api.InputDataFramePtr(api.DataFramePtr("df-in", 2), 123456789012), api.DataFramePtr("df-out", 2), api.CryptoKeyPtr(1), c)
        assert stub.map(ws.task_to_proto(mt), timeout=30) == wp.Empty()
        got = wg.tasks[-1]
        assert isinstance(got, api.MapTask) and got.input_data_frame_ptr.row_count == 123456789012 and got.output_data_frame_ptr.partition == 2
        assert [n[:6] for n in got.circuit.nodes] == [n[:6] for n in c.nodes]
        rc = cm.reduce_add_circuit(8)
        rt = api.ReduceTask("s-1", [api.InputDataFramePtr(api.DataFramePtr("h", k), 3) for k in range(3)], api.DataFramePtr("o", 1), api.CryptoKeyPtr(1), rc)
        assert stub.reduce(ws.task_to_proto(rt), timeout=30) == wp.Empty()
        got = wg.tasks[-1]
        assert isinstance(got, api.ReduceTask) and [p.pointer.partition for p in got.input_data_frame_ptrs] == [0, 1, 2] and got.output_data_frame_ptr.partition == 1
        bad = api.MapTask("bad", mt.input_data_frame_ptr, mt.output_data_frame_ptr, api.CryptoKeyPtr(1), c)
        with pytest.raises(grpc.RpcError) as e:           # non-OK => the client reports Status::ERROR (grpc_worker_group.cpp:112-130)
            stub.map(ws.task_to_proto(bad), timeout=30)
        assert e.value.code() == grpc.StatusCode.INTERNAL and "recorded failure" in e.value.details()
        broken = ws.task_to_proto(mt); broken.circuit.edges[0].end = 10 ** 6
        with pytest.raises(grpc.RpcError) as e:
            stub.map(broken, timeout=30)
        assert e.value.code() == grpc.StatusCode.INVALID_ARGUMENT
        # a 32-bit multiplier (10^4 nodes) fits the 32 MiB caps with room to spare
        big = ws.task_to_proto(api.MapTask("s-1", mt.input_data_frame_ptr, mt.output_data_frame_ptr, api.CryptoKeyPtr(1), cm.multiply_circuit(32)))
        assert big.ByteSize() < ws.MAX_MESSAGE // 8 and stub.map(big, timeout=60) == wp.Empty()
    finally:
        stub.close()
        server.stop(0)
