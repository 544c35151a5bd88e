"""Wire schema of herdsman's worker seam, built at import time without protoc.

Field-for-field mirror of the reference's proto3 files (package `herd.proto`):
  common.proto   Empty, Edge{start=1,end=2}, DataBlob, enum SchemaType{BINFHE=0}, enum DataType{BIT..INT64}
  node.proto     InputNode{tuple_index=1,field_index=2,bit_index=3}, OutputNode{field_index=1,bit_index=2},
                 ConstantNode{value=1}, enum Operation{AND=0,OR=1,NOT=2}, OperationNode{type=1},
                 Node{oneof node{input=1,output=2,constant=3,operation=4}}            (proto/src/node.proto:1-45)
  circuit.proto  InputStructure{fields=1}, OutputColumn{data_type=1,name=2},
                 Circuit{inputs=1,output=2,nodes=3,edges=4}                          (proto/src/circuit.proto)
  worker.proto   DataFramePtr{data_frame_uuid=1,partition=2}, InputDataFramePtr{pointer=1,row_count=4},
                 CryptoKeyPtr{schema_type=1}, MapTask{1..5}, ReduceTask{1..5},
                 service Worker{rpc map(MapTask) returns (Empty); rpc reduce(ReduceTask) returns (Empty)}
                                                                                      (proto/src/worker.proto:11-61)
so that the bytes `herd::mapper::to_proto(MapTask)` produces (mappers/src/worker.cpp:84-124) parse here and vice
versa.  The image has grpcio and protobuf but no protoc / grpc_tools, so the FileDescriptorProtos are written out by
hand and message classes come from the runtime's message factory.
"""
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

_F = descriptor_pb2.FieldDescriptorProto
PACKAGE = "herd.proto"


def _field(msg, name, number, ftype, label=_F.LABEL_OPTIONAL, type_name=None, oneof=None):
    f = msg.field.add()
    f.name, f.number, f.type, f.label = name, number, ftype, label
    if type_name:
        f.type_name = "." + PACKAGE + "." + type_name
    if oneof is not None:
        f.oneof_index = oneof
    return f


def _enum(fd, name, values):
    e = fd.enum_type.add()
    e.name = name
    for k, v in values:
        x = e.value.add(); x.name, x.number = k, v


def _common():
    fd = descriptor_pb2.FileDescriptorProto(name="common.proto", package=PACKAGE, syntax="proto3")
    fd.message_type.add().name = "Empty"
    m = fd.message_type.add(); m.name = "Edge"
    _field(m, "start", 1, _F.TYPE_UINT32); _field(m, "end", 2, _F.TYPE_UINT32)
    m = fd.message_type.add(); m.name = "DataBlob"
    _field(m, "blob", 1, _F.TYPE_BYTES)
    _enum(fd, "SchemaType", [("BINFHE", 0)])
    _enum(fd, "DataType", [("BIT", 0), ("UINT8", 1), ("UINT16", 2), ("UINT32", 3), ("UINT64", 4),
                           ("INT8", 5), ("INT16", 6), ("INT32", 7), ("INT64", 8)])
    return fd


def _node():
    fd = descriptor_pb2.FileDescriptorProto(name="node.proto", package=PACKAGE, syntax="proto3")
    m = fd.message_type.add(); m.name = "InputNode"
    _field(m, "tuple_index", 1, _F.TYPE_UINT32); _field(m, "field_index", 2, _F.TYPE_UINT32); _field(m, "bit_index", 3, _F.TYPE_UINT32)
    m = fd.message_type.add(); m.name = "OutputNode"
    _field(m, "field_index", 1, _F.TYPE_UINT32); _field(m, "bit_index", 2, _F.TYPE_UINT32)
    m = fd.message_type.add(); m.name = "ConstantNode"
    _field(m, "value", 1, _F.TYPE_BOOL)
    _enum(fd, "Operation", [("AND", 0), ("OR", 1), ("NOT", 2)])
    m = fd.message_type.add(); m.name = "OperationNode"
    _field(m, "type", 1, _F.TYPE_ENUM, type_name="Operation")
    m = fd.message_type.add(); m.name = "Node"
    m.oneof_decl.add().name = "node"
    _field(m, "input", 1, _F.TYPE_MESSAGE, type_name="InputNode", oneof=0)
    _field(m, "output", 2, _F.TYPE_MESSAGE, type_name="OutputNode", oneof=0)
    _field(m, "constant", 3, _F.TYPE_MESSAGE, type_name="ConstantNode", oneof=0)
    _field(m, "operation", 4, _F.TYPE_MESSAGE, type_name="OperationNode", oneof=0)
    return fd


def _circuit():
    fd = descriptor_pb2.FileDescriptorProto(name="circuit.proto", package=PACKAGE, syntax="proto3",
                                            dependency=["common.proto", "node.proto"])
    m = fd.message_type.add(); m.name = "InputStructure"
    _field(m, "fields", 1, _F.TYPE_ENUM, _F.LABEL_REPEATED, "DataType")
    m = fd.message_type.add(); m.name = "OutputColumn"
    _field(m, "data_type", 1, _F.TYPE_ENUM, type_name="DataType"); _field(m, "name", 2, _F.TYPE_STRING)
    m = fd.message_type.add(); m.name = "Circuit"
    _field(m, "inputs", 1, _F.TYPE_MESSAGE, _F.LABEL_REPEATED, "InputStructure")
    _field(m, "output", 2, _F.TYPE_MESSAGE, _F.LABEL_REPEATED, "OutputColumn")
    _field(m, "nodes", 3, _F.TYPE_MESSAGE, _F.LABEL_REPEATED, "Node")
    _field(m, "edges", 4, _F.TYPE_MESSAGE, _F.LABEL_REPEATED, "Edge")
    return fd


def _worker():
    fd = descriptor_pb2.FileDescriptorProto(name="worker.proto", package=PACKAGE, syntax="proto3",
                                            dependency=["common.proto", "circuit.proto"])
    m = fd.message_type.add(); m.name = "DataFramePtr"
    _field(m, "data_frame_uuid", 1, _F.TYPE_STRING); _field(m, "partition", 2, _F.TYPE_UINT32)
    m = fd.message_type.add(); m.name = "InputDataFramePtr"
    _field(m, "pointer", 1, _F.TYPE_MESSAGE, type_name="DataFramePtr"); _field(m, "row_count", 4, _F.TYPE_UINT64)
    m = fd.message_type.add(); m.name = "CryptoKeyPtr"
    _field(m, "schema_type", 1, _F.TYPE_ENUM, type_name="SchemaType")
    for name, rep in (("MapTask", False), ("ReduceTask", True)):
        m = fd.message_type.add(); m.name = name
        _field(m, "session_uuid", 1, _F.TYPE_STRING)
        if rep:
            _field(m, "input_data_frame_ptrs", 2, _F.TYPE_MESSAGE, _F.LABEL_REPEATED, "InputDataFramePtr")
        else:
            _field(m, "input_data_frame_ptr", 2, _F.TYPE_MESSAGE, type_name="InputDataFramePtr")
        _field(m, "output_data_frame_ptr", 3, _F.TYPE_MESSAGE, type_name="DataFramePtr")
        _field(m, "crypto_key_ptr", 4, _F.TYPE_MESSAGE, type_name="CryptoKeyPtr")
        _field(m, "circuit", 5, _F.TYPE_MESSAGE, type_name="Circuit")
    _enum(fd, "CompletionStatus", [("SUCCESS", 0), ("TIMED_OUT", 1), ("ERROR", 2)])
    svc = fd.service.add(); svc.name = "Worker"
    for rpc, req in (("map", "MapTask"), ("reduce", "ReduceTask")):
        mth = svc.method.add()
        mth.name, mth.input_type, mth.output_type = rpc, "." + PACKAGE + "." + req, "." + PACKAGE + ".Empty"
    return fd


_POOL = descriptor_pool.DescriptorPool()
FILES = [_common(), _node(), _circuit(), _worker()]
for _fd in FILES:
    _POOL.Add(_fd)


def _cls(name):
    return message_factory.GetMessageClass(_POOL.FindMessageTypeByName(PACKAGE + "." + name))


Empty, Edge = _cls("Empty"), _cls("Edge")
InputNode, OutputNode, ConstantNode, OperationNode, Node = (_cls(n) for n in ("InputNode", "OutputNode", "ConstantNode", "OperationNode", "Node"))
InputStructure, OutputColumn, Circuit = _cls("InputStructure"), _cls("OutputColumn"), _cls("Circuit")
DataFramePtr, InputDataFramePtr, CryptoKeyPtr = _cls("DataFramePtr"), _cls("InputDataFramePtr"), _cls("CryptoKeyPtr")
MapTask, ReduceTask = _cls("MapTask"), _cls("ReduceTask")

SERVICE = PACKAGE + ".Worker"                    # method paths /herd.proto.Worker/map and /herd.proto.Worker/reduce
# proto::SchemaType -> int(herd::common::SchemaType) (mappers/src/crypto.cpp:7-16; schema_type.hpp:9-13: NONE = 0, BINFHE = 1)
SCHEMA_TO_MODEL = {0: 1}
MODEL_TO_SCHEMA = {1: 0}
