"""gRPC `Worker` façade over the B200 backend: stock herdsman, configured with
`workers: { grpc: { addresses: [ host:port ] } }` (template.yaml:10-14), drives this process exactly as it drives
the CPU herd worker -- zero herdsman changes.

  service Worker { rpc map(MapTask) returns (Empty); rpc reduce(ReduceTask) returns (Empty); }
                                                                      proto/src/worker.proto:57-61
  client side: GrpcWorkerGroup::schedule_task serialises herd::mapper::to_proto(task) and calls Asyncmap /
  Asyncreduce on an insecure channel with 32 MiB message caps (grpc_worker_group.cpp:13-26,70-105); the RPC
  returns when the partition is done; status OK => COMPLETED, anything else => ERROR (grpc_worker_group.cpp:112-130).

The handlers rebuild the circuit the way herd::mapper::to_model(proto::Circuit) does (mappers/src/executor.cpp:81-121:
nodes inserted in order so node id == index, edges added in message order so operand order == edge order) and hand
the task to the C ABI through the same CudaWorkerGroup mirror the in-process adapter uses.  Host side in Python
by necessity: the image has grpcio + protobuf but no protoc / gRPC C++ (message classes come from hand-built
descriptors, worker_proto.py).

Run:  python -m herdsman_b200.worker_server --storage-dir DIR --key-dir DIR [--port 5001] [--devices 0,1,...]
The worker shares herdsman's storage and key directories (the reference's workers do: the Lambda group watches the
storage directory for the worker's output file, lambda_http_worker_group.cpp:69-72,304).
"""
import argparse
from concurrent import futures

import grpc

from . import api, worker_proto as wp
from .circuit import Circuit, DataType, NodeKind, Operation

MAX_MESSAGE = 32 * 1024 * 1024                      # grpc_worker_group.cpp:18-20, main.cpp:135-136
TASK_TIMEOUT_S = 24 * 3600


class MappingError(ValueError):
    """herd::mapper::MappingError: the message does not map onto the model"""


# ------------------------------------------------------------------------------------------------ proto -> model
def circuit_to_model(pb) -> Circuit:
    """herd::mapper::to_model(const proto::Circuit&)  (mappers/src/executor.cpp:81-121)"""
    try:
        c = Circuit([[DataType(t) for t in s.fields] for s in pb.inputs], [(col.name, DataType(col.data_type)) for col in pb.output])
    except ValueError as e:
        raise MappingError("Proto schema, model mismatch") from e
    for node in pb.nodes:                             # graph.insert(to_model(node)): node id = position in the message
        kind = node.WhichOneof("node")
        if kind == "input":
            c._add(NodeKind.INPUT, tup=node.input.tuple_index, field=node.input.field_index, bit=node.input.bit_index)
        elif kind == "output":
            c._add(NodeKind.OUTPUT, field=node.output.field_index, bit=node.output.bit_index)
        elif kind == "constant":
            c._add(NodeKind.CONSTANT, const=1 if node.constant.value else 0)
        elif kind == "operation":
            if node.operation.type not in (0, 1, 2):  # wire enum has AND / OR / NOT only (node.proto:25-30)
                raise MappingError("Proto schema, model mismatch")
            c._add(NodeKind.OPERATION, op=Operation(node.operation.type))
        else:
            raise MappingError("Proto schema, model mismatch")
    n = len(c.nodes)
    for e in pb.edges:                                # graph.add_edge(begin + start, begin + end), in message order
        if e.start >= n or e.end >= n:
            raise MappingError("edge out of range")
        c.edges.append((int(e.start), int(e.end)))
    return c


def _frame_ptr(pb):
    if not pb.data_frame_uuid:
        raise MappingError("missing data frame uuid")
    return api.DataFramePtr(pb.data_frame_uuid, pb.partition)


def _schema(pb):
    try:
        return api.CryptoKeyPtr(wp.SCHEMA_TO_MODEL[pb.schema_type])      # proto BINFHE (0) -> SchemaType::BINFHE (1) -> "1.key"
    except KeyError as e:
        raise MappingError("Proto schema, model mismatch") from e


def map_task_to_model(pb) -> api.MapTask:
    return api.MapTask(pb.session_uuid,
                       api.InputDataFramePtr(_frame_ptr(pb.input_data_frame_ptr.pointer), pb.input_data_frame_ptr.row_count),
                       _frame_ptr(pb.output_data_frame_ptr), _schema(pb.crypto_key_ptr), circuit_to_model(pb.circuit))


def reduce_task_to_model(pb) -> api.ReduceTask:
    return api.ReduceTask(pb.session_uuid,
                          [api.InputDataFramePtr(_frame_ptr(p.pointer), p.row_count) for p in pb.input_data_frame_ptrs],
                          _frame_ptr(pb.output_data_frame_ptr), _schema(pb.crypto_key_ptr), circuit_to_model(pb.circuit))


# ------------------------------------------------------------------------------------------------ model -> proto
# (what herdsman's side does, mappers/src/worker.cpp:84-124 and executor.cpp:231-336; used by the tests as the client)
def circuit_to_proto(c: Circuit):
    pb = wp.Circuit()
    for tup in c.inputs:
        pb.inputs.add().fields.extend(int(t) for t in tup)
    for name, t in c.output:
        col = pb.output.add(); col.name = name; col.data_type = int(t)
    for kind, op, const, tup, field, bit, _flags in c.nodes:
        node = pb.nodes.add()
        if kind == NodeKind.INPUT:
            node.input.tuple_index, node.input.field_index, node.input.bit_index = tup, field, bit
            node.input.SetInParent()
        elif kind == NodeKind.OUTPUT:
            node.output.field_index, node.output.bit_index = field, bit
            node.output.SetInParent()
        elif kind == NodeKind.CONSTANT:
            node.constant.value = bool(const)
            node.constant.SetInParent()
        else:
            if op not in (Operation.AND, Operation.OR, Operation.NOT):
                raise MappingError("Not supported")           # to_proto(Operation) throws for XOR / NAND / MUX (executor.cpp:242-245)
            node.operation.type = int(op)
            node.operation.SetInParent()
    for start, end in sorted(c.edges, key=lambda e: e[0]):        # DAG::edges() is a multimap keyed by the start node (dag.hpp:453-475)
        e = pb.edges.add(); e.start, e.end = start, end
    return pb


def _ptr_to_proto(pb, p):
    pb.data_frame_uuid, pb.partition = p.uuid, p.partition


def task_to_proto(task):
    if isinstance(task, api.ReduceTask):
        pb = wp.ReduceTask()
        for p in task.input_data_frame_ptrs:
            q = pb.input_data_frame_ptrs.add(); _ptr_to_proto(q.pointer, p.pointer); q.row_count = p.row_count
    else:
        pb = wp.MapTask()
        _ptr_to_proto(pb.input_data_frame_ptr.pointer, task.input_data_frame_ptr.pointer)
        pb.input_data_frame_ptr.row_count = task.input_data_frame_ptr.row_count
    pb.session_uuid = task.session_uuid
    _ptr_to_proto(pb.output_data_frame_ptr, task.output_data_frame_ptr)
    pb.crypto_key_ptr.schema_type = wp.MODEL_TO_SCHEMA[int(task.crypto_key_ptr.schema_type)]
    pb.crypto_key_ptr.SetInParent()
    pb.circuit.CopyFrom(circuit_to_proto(task.circuit))
    return pb


# ------------------------------------------------------------------------------------------------ the service
class WorkerService:
    """implements herd.proto.Worker on top of CudaWorkerGroup (IWorkerGroup mirror, api.py)"""

    def __init__(self, worker_group):
        self.wg = worker_group

    def _run(self, to_model, request, context):
        try:
            task = to_model(request)
        except MappingError as e:
            context.abort(grpc.StatusCode.INVALID_ARGUMENT, str(e))
        status = self.wg.schedule_task(task).wait(TASK_TIMEOUT_S)
        if status != api.TaskHandle.Status.COMPLETED:
            # the client maps every non-OK status to Status::ERROR (grpc_worker_group.cpp:112-130)
            detail = self.wg.backend.L.herdb_last_error(self.wg.backend.h).decode()
            code = grpc.StatusCode.RESOURCE_EXHAUSTED if status == api.TaskHandle.Status.TIME_OUT else grpc.StatusCode.INTERNAL
            context.abort(code, "task failed: " + detail)
        return wp.Empty()

    def map(self, request, context):
        return self._run(map_task_to_model, request, context)

    def reduce(self, request, context):
        return self._run(reduce_task_to_model, request, context)

    def handler(self):
        return grpc.method_handlers_generic_handler(wp.SERVICE, {
            "map": grpc.unary_unary_rpc_method_handler(self.map, request_deserializer=wp.MapTask.FromString,
                                                       response_serializer=lambda m: m.SerializeToString()),
            "reduce": grpc.unary_unary_rpc_method_handler(self.reduce, request_deserializer=wp.ReduceTask.FromString,
                                                          response_serializer=lambda m: m.SerializeToString()),
        })


def serve(worker_group, address="127.0.0.1:0", max_workers=None):
    """starts the server; returns (grpc.Server, bound port).  One RPC thread per task the backend runs at once, plus
    a few so that refusals are answered while long tasks run."""
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers or worker_group.concurrent_workers() + 4),
                         options=[("grpc.max_send_message_length", MAX_MESSAGE), ("grpc.max_receive_message_length", MAX_MESSAGE)])
    server.add_generic_rpc_handlers((WorkerService(worker_group).handler(),))
    port = server.add_insecure_port(address)           # the reference's channel is insecure (grpc_worker_group.cpp:24)
    if not port:
        raise RuntimeError("cannot bind " + address)
    server.start()
    return server, port


class WorkerStub:
    """client side of the seam as GrpcWorkerGroup uses it: Worker::NewStub(channel)->map / reduce"""

    def __init__(self, target):
        self.channel = grpc.insecure_channel(target, options=[("grpc.max_send_message_length", MAX_MESSAGE),
                                                              ("grpc.max_receive_message_length", MAX_MESSAGE)])
        ser = lambda m: m.SerializeToString()           # noqa: E731
        self.map = self.channel.unary_unary("/" + wp.SERVICE + "/map", request_serializer=ser, response_deserializer=wp.Empty.FromString)
        self.reduce = self.channel.unary_unary("/" + wp.SERVICE + "/reduce", request_serializer=ser, response_deserializer=wp.Empty.FromString)

    def close(self):
        self.channel.close()


def main():
    ap = argparse.ArgumentParser(description="herd.proto.Worker served by the B200 backend")
    ap.add_argument("--storage-dir", required=True, help="herdsman's server.storage_directory")
    ap.add_argument("--key-dir", required=True, help="herdsman's server.key_directory")
    ap.add_argument("--host", default="0.0.0.0")
    ap.add_argument("--port", type=int, default=5001)
    ap.add_argument("--devices", default="0", help="comma-separated CUDA device ids")
    args = ap.parse_args()
    wg = api.CudaWorkerGroup(args.storage_dir, args.key_dir, devices=tuple(int(d) for d in args.devices.split(",")))
    server, port = serve(wg, "%s:%d" % (args.host, args.port))
    print("herd.proto.Worker on port %d, devices %s" % (port, args.devices), flush=True)
    try:
        server.wait_for_termination()
    finally:
        wg.close()


if __name__ == "__main__":
    main()
