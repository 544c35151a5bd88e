"""Host mirror of herdsman's circuit model, plus the circuit builders the benchmarks use.

Mirrors (same names, same meaning) the reference types
  herd::common::Circuit      common/include/herd/common/model/circuit/circuit.hpp:14-30
  node_t = Input/Output/Constant/Operation  .../circuit/nodes.hpp:12-43
  Operation {AND, OR, NOT, XOR, NAND, MUX}  .../circuit/operation.hpp:12-20
  DataType                   common/include/herd/common/model/data_type.hpp:9-22
  DAG<T>: operands of a node = parents() in edge-insertion order   common/include/herd/common/dag.hpp:208-230
and flattens them into the `herdb_circuit` struct of include/herdb.h for the C ABI.
"""
import ctypes as C
import enum

import numpy as np


class DataType(enum.IntEnum):
    BIT = 0
    UINT8 = 1
    UINT16 = 2
    UINT32 = 3
    UINT64 = 4
    INT8 = 5
    INT16 = 6
    INT32 = 7
    INT64 = 8


BIT_WIDTH = {DataType.BIT: 1, DataType.UINT8: 8, DataType.UINT16: 16, DataType.UINT32: 32, DataType.UINT64: 64,
             DataType.INT8: 8, DataType.INT16: 16, DataType.INT32: 32, DataType.INT64: 64}


class Operation(enum.IntEnum):
    AND = 0
    OR = 1
    NOT = 2
    XOR = 3
    NAND = 4
    MUX = 5


class NodeKind(enum.IntEnum):
    INPUT = 0
    OUTPUT = 1
    CONSTANT = 2
    OPERATION = 3


class herdb_node(C.Structure):
    _fields_ = [("kind", C.c_uint8), ("op", C.c_uint8), ("const_val", C.c_uint8), ("flags", C.c_uint8),
                ("tuple", C.c_uint32), ("field", C.c_uint32), ("bit", C.c_uint32)]


class herdb_edge(C.Structure):
    _fields_ = [("from_", C.c_uint32), ("to", C.c_uint32)]


class herdb_circuit(C.Structure):
    _fields_ = [("nodes", C.POINTER(herdb_node)), ("n_nodes", C.c_size_t),
                ("edges", C.POINTER(herdb_edge)), ("n_edges", C.c_size_t),
                ("in_field_types", C.POINTER(C.c_uint8)), ("in_tuple_offsets", C.POINTER(C.c_uint32)),
                ("n_tuples", C.c_size_t),
                ("out_field_types", C.POINTER(C.c_uint8)), ("n_out_fields", C.c_size_t)]


class herdb_circuit_info(C.Structure):
    _fields_ = [("in_bits", C.c_uint32), ("out_bits", C.c_uint32), ("boot_gates", C.c_uint32),
                ("levels", C.c_uint32), ("max_level_width", C.c_uint32), ("slots", C.c_uint32)]


class herdb_schedule_info(C.Structure):
    _fields_ = [("levels", C.c_uint32), ("max_level_width", C.c_uint32), ("slots", C.c_uint32), ("policy", C.c_uint32),
                ("rounds_asap", C.c_uint64), ("rounds", C.c_uint64)]


class Circuit:
    """herd::common::Circuit: inputs[tuple][field] -> DataType, output[] -> (name, DataType), circuit_graph."""

    def __init__(self, inputs, output):
        self.inputs = [[DataType(t) for t in tup] for tup in inputs]
        self.output = [(name, DataType(t)) for name, t in output]
        self.nodes = []       # (kind, op, const, tuple, field, bit)
        self.edges = []       # (from, to) in insertion order
        self._keep = None

    # -- graph construction (DAG::emplace / add_edge) ------------------------------------------
    def _add(self, kind, op=0, const=0, tup=0, field=0, bit=0, parents=(), flags=0):
        self.nodes.append((int(kind), int(op), int(const), int(tup), int(field), int(bit), int(flags)))
        nid = len(self.nodes) - 1
        for p in parents:
            self.edges.append((int(p), nid))
        return nid

    def input(self, tup, field, bit):
        return self._add(NodeKind.INPUT, tup=tup, field=field, bit=bit)

    def constant(self, value):
        return self._add(NodeKind.CONSTANT, const=1 if value else 0)

    def op(self, operation, *parents, flags=0):
        return self._add(NodeKind.OPERATION, op=Operation(operation), parents=parents, flags=flags)

    def output_bit(self, field, bit, parent):
        return self._add(NodeKind.OUTPUT, field=field, bit=bit, parents=(parent,))

    def and_(self, a, b): return self.op(Operation.AND, a, b)
    def or_(self, a, b): return self.op(Operation.OR, a, b)
    def not_(self, a): return self.op(Operation.NOT, a)
    def nand(self, a, b): return self.op(Operation.NAND, a, b)
    def xor(self, a, b): return self.op(Operation.XOR, a, b)
    def xor_fast(self, a, b):
        """one bootstrap (OpenFHE BINGATE XOR_FAST) instead of the three of EvalBinGate(XOR): HERDB_NODE_FAST"""
        return self.op(Operation.XOR, a, b, flags=1)
    def mux(self, sel, a, b): return self.op(Operation.MUX, sel, a, b)

    def xor_wire(self, a, b):
        """XOR in the AND/OR/NOT basis the herd wire format can carry (node.proto:25-30;
        mappers/src/executor.cpp:231-249 rejects XOR/NAND/MUX): OR(AND(a,!b), AND(!a,b))."""
        return self.or_(self.and_(a, self.not_(b)), self.and_(self.not_(a), b))

    def input_word(self, tup, field):
        return [self.input(tup, field, b) for b in range(BIT_WIDTH[self.inputs[tup][field]])]

    def output_word(self, field, bits):
        for b, w in enumerate(bits):
            self.output_bit(field, b, w)

    # -- facts -----------------------------------------------------------------------------------
    @property
    def in_bits(self):
        return sum(BIT_WIDTH[t] for tup in self.inputs for t in tup)

    @property
    def out_bits(self):
        return sum(BIT_WIDTH[t] for _, t in self.output)

    # -- C ABI -----------------------------------------------------------------------------------
    def to_c(self):
        nodes = (herdb_node * max(1, len(self.nodes)))()
        for i, (k, op, cv, tp, fd, bt, fl) in enumerate(self.nodes):
            nodes[i] = herdb_node(k, op, cv, fl, tp, fd, bt)
        edges = (herdb_edge * max(1, len(self.edges)))()
        for i, (f, t) in enumerate(self.edges):
            edges[i] = herdb_edge(f, t)
        flat = [int(t) for tup in self.inputs for t in tup]
        in_types = (C.c_uint8 * max(1, len(flat)))(*flat)
        offs = [0]
        for tup in self.inputs:
            offs.append(offs[-1] + len(tup))
        in_offs = (C.c_uint32 * len(offs))(*offs)
        out_types = (C.c_uint8 * max(1, len(self.output)))(*[int(t) for _, t in self.output])
        cc = herdb_circuit(nodes, len(self.nodes), edges, len(self.edges), in_types, in_offs, len(self.inputs),
                           out_types, len(self.output))
        self._keep = (nodes, edges, in_types, in_offs, out_types)      # keep the buffers alive
        return cc


# ------------------------------------------------------------------------------------------------
# circuit builders for the BASELINE.json configs.  `basis="wire"` uses only AND/OR/NOT (what a herd
# client can actually send); `basis="model"` uses the XOR / MUX operations of operation.hpp.
def _xor(c, a, b, basis):
    """basis "wire": AND/OR/NOT only (what herdsman's proto can carry); "model": XOR nodes, lowered by the backend
    to the three-gate form of v1.0.x EvalBinGate(XOR); "fast": XOR nodes flagged HERDB_NODE_FAST, one bootstrap each"""
    return c.xor_wire(a, b) if basis == "wire" else c.xor_fast(a, b) if basis == "fast" else c.xor(a, b)


def _full_adder(c, a, b, cin, basis):
    axb = _xor(c, a, b, basis)
    s = _xor(c, axb, cin, basis)
    cout = c.or_(c.and_(a, b), c.and_(axb, cin))
    return s, cout


def _ripple_add(c, xs, ys, basis, cin=None, want_carry=False):
    out = []
    carry = cin
    for i, (a, b) in enumerate(zip(xs, ys)):
        last = i == len(xs) - 1
        if carry is None:
            out.append(_xor(c, a, b, basis))
            if not last or want_carry:
                carry = c.and_(a, b)
        elif last and not want_carry:
            out.append(_xor(c, _xor(c, a, b, basis), carry, basis))
        else:
            s, carry = _full_adder(c, a, b, carry, basis)
            out.append(s)
    return (out, carry) if want_carry else out


def _less_than(c, xs, ys, basis):
    """unsigned x < y, LSB-first ripple: lt = (!x_i & y_i) | (!(x_i ^ y_i) & lt)"""
    lt = None
    for a, b in zip(xs, ys):
        nb = c.and_(c.not_(a), b)
        if lt is None:
            lt = nb
        else:
            eq = c.not_(_xor(c, a, b, basis))
            lt = c.or_(nb, c.and_(eq, lt))
    return lt


def adder_circuit(bits=8, basis="wire"):
    """config 1: two UINT<bits> columns -> one UINT<bits> column, (x + y) mod 2^bits."""
    t = {8: DataType.UINT8, 16: DataType.UINT16, 32: DataType.UINT32, 64: DataType.UINT64}[bits]
    c = Circuit([[t, t]], [("sum", t)])
    c.output_word(0, _ripple_add(c, c.input_word(0, 0), c.input_word(0, 1), basis))
    return c


def compare_circuit(bits=32, basis="wire"):
    """config 3: two UINT<bits> columns -> one BIT column, x < y."""
    t = {8: DataType.UINT8, 16: DataType.UINT16, 32: DataType.UINT32, 64: DataType.UINT64}[bits]
    c = Circuit([[t, t]], [("lt", DataType.BIT)])
    c.output_bit(0, 0, _less_than(c, c.input_word(0, 0), c.input_word(0, 1), basis))
    return c


def multiply_circuit(bits=32, basis="wire"):
    """config 4: two UINT<bits> columns -> low <bits> bits of x*y (shift-and-add array multiplier)."""
    t = {8: DataType.UINT8, 16: DataType.UINT16, 32: DataType.UINT32, 64: DataType.UINT64}[bits]
    c = Circuit([[t, t]], [("prod", t)])
    xs, ys = c.input_word(0, 0), c.input_word(0, 1)
    acc = [c.and_(xs[j], ys[0]) for j in range(bits)]
    for i in range(1, bits):
        pp = [c.and_(xs[j], ys[i]) for j in range(bits - i)]
        acc = acc[:i] + _ripple_add(c, acc[i:], pp, basis)
    c.output_word(0, acc)
    return c


def select_circuit(bits=8, basis="wire"):
    """config 5, last stage: (BIT sel, UINT a, UINT b) -> sel ? a : b."""
    t = {8: DataType.UINT8, 16: DataType.UINT16, 32: DataType.UINT32, 64: DataType.UINT64}[bits]
    c = Circuit([[DataType.BIT, t, t]], [("sel", t)])
    s = c.input(0, 0, 0)
    xs, ys = c.input_word(0, 1), c.input_word(0, 2)
    out = []
    for a, b in zip(xs, ys):
        out.append(c.mux(s, a, b) if basis == "model" else c.or_(c.and_(s, a), c.and_(c.not_(s), b)))
    c.output_word(0, out)
    return c


def pipeline_stages(bits=8, basis="wire"):
    """config 5 as three chained map stages whose rows feed one another unchanged (herdb_eval_pipeline_rows):
    (a, b, c) -> (a + b, c) -> (a + b < c, a + b, c) -> select.  Columns a later stage needs ride along as
    pass-through outputs (no bootstrap)."""
    t = {8: DataType.UINT8, 16: DataType.UINT16, 32: DataType.UINT32}[bits]
    s1 = Circuit([[t, t, t]], [("s", t), ("c", t)])
    s1.output_word(0, _ripple_add(s1, s1.input_word(0, 0), s1.input_word(0, 1), basis)); s1.output_word(1, s1.input_word(0, 2))
    s2 = Circuit([[t, t]], [("lt", DataType.BIT), ("s", t), ("c", t)])
    xs, ys = s2.input_word(0, 0), s2.input_word(0, 1)
    s2.output_bit(0, 0, _less_than(s2, xs, ys, basis)); s2.output_word(1, xs); s2.output_word(2, ys)
    return [s1, s2, select_circuit(bits, basis)]


def pipeline_circuit(bits=8, basis="wire"):
    """config 5 fused into one map stage: (a, b, c) -> (a + b < c) ? a + b : c."""
    t = {8: DataType.UINT8, 16: DataType.UINT16, 32: DataType.UINT32}[bits]
    c = Circuit([[t, t, t]], [("out", t)])
    a, b, cc = c.input_word(0, 0), c.input_word(0, 1), c.input_word(0, 2)
    s = _ripple_add(c, a, b, basis)
    lt = _less_than(c, s, cc, basis)
    out = [c.or_(c.and_(lt, x), c.and_(c.not_(lt), y)) for x, y in zip(s, cc)]
    c.output_word(0, out)
    return c


def reduce_add_circuit(bits=8, basis="wire"):
    """binary reducer of a ReduceStage (stage.hpp:29-34): two rows of one UINT<bits> column -> their sum mod 2^bits"""
    t = {8: DataType.UINT8, 16: DataType.UINT16, 32: DataType.UINT32, 64: DataType.UINT64}[bits]
    c = Circuit([[t], [t]], [("sum", t)])
    c.output_word(0, _ripple_add(c, c.input_word(0, 0), c.input_word(1, 0), basis))
    return c


def reduce_max_circuit(bits=8, basis="wire"):
    """binary reducer: two rows of one UINT<bits> column -> the larger value"""
    t = {8: DataType.UINT8, 16: DataType.UINT16, 32: DataType.UINT32, 64: DataType.UINT64}[bits]
    c = Circuit([[t], [t]], [("max", t)])
    xs, ys = c.input_word(0, 0), c.input_word(1, 0)
    lt = _less_than(c, xs, ys, basis)
    c.output_word(0, [c.or_(c.and_(lt, y), c.and_(c.not_(lt), x)) for x, y in zip(xs, ys)])
    return c


def single_gate_circuit(op=Operation.NAND):
    """config 2 as a circuit: (BIT, BIT) -> BIT."""
    c = Circuit([[DataType.BIT, DataType.BIT]], [("o", DataType.BIT)])
    c.output_bit(0, 0, c.op(op, c.input(0, 0, 0), c.input(0, 1, 0)))
    return c


def pack_words(values, bits):
    """integers -> LSB-first bit matrix [len, bits] (uint8)"""
    v = np.asarray(values, dtype=np.uint64).reshape(-1, 1)
    return ((v >> np.arange(bits, dtype=np.uint64)) & 1).astype(np.uint8)


def unpack_words(bitmat):
    b = np.asarray(bitmat, dtype=np.uint64)
    return (b << np.arange(b.shape[1], dtype=np.uint64)).sum(axis=1)
