// circuit.cpp -- see circuit.h
#include "circuit.h"

#include <algorithm>
#include <queue>

namespace herdb {

namespace {

const u32 kTypeBits[9] = {1, 8, 16, 32, 64, 8, 16, 32, 64};   // native_type_mapping.cpp:8-35

struct Value {                    // what a node evaluates to, without materialising NOT / constants
    uint8_t pool = kPoolIn;       // kPoolIn: input bit idx;  kPoolSlots: boot-gate id (slot assigned later)
    uint8_t flags = 0;            // OperandFlags
    u32 idx = 0;
    u32 level = 0;                // 0 for inputs / constants, else level of the producing gate (1-based)
};

struct BootGate {
    uint8_t gate;
    Value a, b;
    u32 level;                    // 1-based
    u32 last_use = 0;             // last level that reads the result (UINT32_MAX if an output reads it)
    u32 slot = 0;
};

Value negate(Value v) {
    if (v.flags & kOpConst) v.flags ^= kOpConstVal; else v.flags ^= kOpNeg;
    return v;
}

}  // namespace

int compile_circuit(const herdb_circuit* c, CompiledCircuit& out, std::string& err) {
    out = CompiledCircuit();
    if (!c || (!c->nodes && c->n_nodes) || (!c->edges && c->n_edges) || !c->in_tuple_offsets) { err = "null circuit"; return HERDB_ERR_ARG; }
    const u32 nn = (u32)c->n_nodes;
    // field offsets
    const size_t n_in_fields = c->in_tuple_offsets[c->n_tuples];
    std::vector<u32> in_off(n_in_fields + 1, 0), out_off(c->n_out_fields + 1, 0);
    for (size_t f = 0; f < n_in_fields; ++f) {
        if (c->in_field_types[f] > HERDB_INT64) { err = "bad input data type"; return HERDB_ERR_ARG; }
        in_off[f + 1] = in_off[f] + kTypeBits[c->in_field_types[f]];
    }
    for (size_t f = 0; f < c->n_out_fields; ++f) {
        if (c->out_field_types[f] > HERDB_INT64) { err = "bad output data type"; return HERDB_ERR_ARG; }
        out_off[f + 1] = out_off[f] + kTypeBits[c->out_field_types[f]];
    }
    out.in_bits = in_off[n_in_fields];
    out.out_bits = out_off[c->n_out_fields];

    // parents in edge order (DAG::parents(), dag.hpp:208-230), children for Kahn
    std::vector<std::vector<u32>> parents(nn), children(nn);
    for (size_t e = 0; e < c->n_edges; ++e) {
        const u32 f = c->edges[e].from, t = c->edges[e].to;
        if (f >= nn || t >= nn) { err = "edge out of range"; return HERDB_ERR_ARG; }
        parents[t].push_back(f); children[f].push_back(t);
    }
    std::vector<u32> indeg(nn), order; order.reserve(nn);
    std::priority_queue<u32, std::vector<u32>, std::greater<u32>> ready;
    for (u32 i = 0; i < nn; ++i) { indeg[i] = (u32)parents[i].size(); if (!indeg[i]) ready.push(i); }
    while (!ready.empty()) {
        u32 i = ready.top(); ready.pop(); order.push_back(i);
        for (u32 ch : children[i]) if (--indeg[ch] == 0) ready.push(ch);
    }
    if (order.size() != nn) { err = "circuit graph has a cycle"; return HERDB_ERR_ARG; }

    std::vector<Value> val(nn);
    std::vector<BootGate> gates;
    std::vector<Value> out_val(out.out_bits);
    std::vector<uint8_t> out_set(out.out_bits, 0);

    auto add_gate = [&](uint8_t code, const Value& a, const Value& b) -> Value {
        BootGate g; g.gate = code; g.a = a; g.b = b; g.level = std::max(a.level, b.level) + 1;
        gates.push_back(g);
        Value v; v.pool = kPoolSlots; v.idx = (u32)gates.size() - 1; v.level = g.level;
        return v;
    };

    for (u32 id : order) {
        const herdb_node& nd = c->nodes[id];
        const auto& par = parents[id];
        switch (nd.kind) {
        case HERDB_NODE_INPUT: {
            if (!par.empty()) { err = "input node with parents"; return HERDB_ERR_ARG; }
            if (nd.tuple >= c->n_tuples) { err = "input tuple out of range"; return HERDB_ERR_ARG; }
            const u32 f = c->in_tuple_offsets[nd.tuple] + nd.field;
            if (f >= c->in_tuple_offsets[nd.tuple + 1] || nd.bit >= kTypeBits[c->in_field_types[f]]) { err = "input field/bit out of range"; return HERDB_ERR_ARG; }
            Value v; v.pool = kPoolIn; v.idx = in_off[f] + nd.bit; val[id] = v;
            break; }
        case HERDB_NODE_CONSTANT: {
            if (!par.empty()) { err = "constant node with parents"; return HERDB_ERR_ARG; }
            Value v; v.flags = kOpConst | (nd.const_val ? kOpConstVal : 0); val[id] = v;
            break; }
        case HERDB_NODE_OUTPUT: {
            if (par.size() != 1) { err = "output node needs exactly one parent"; return HERDB_ERR_ARG; }
            if (nd.field >= c->n_out_fields || nd.bit >= kTypeBits[c->out_field_types[nd.field]]) { err = "output field/bit out of range"; return HERDB_ERR_ARG; }
            const u32 o = out_off[nd.field] + nd.bit;
            out_val[o] = val[par[0]]; out_set[o] = 1;
            val[id] = val[par[0]];
            break; }
        case HERDB_NODE_OPERATION: {
            const size_t want = nd.op == HERDB_OP_NOT ? 1 : nd.op == HERDB_OP_MUX ? 3 : 2;
            if (nd.op > HERDB_OP_MUX) { err = "unknown operation"; return HERDB_ERR_ARG; }
            if (par.size() != want) { err = "operation arity mismatch"; return HERDB_ERR_ARG; }
            const Value a = val[par[0]];
            switch (nd.op) {
            case HERDB_OP_NOT:  val[id] = negate(a); break;
            case HERDB_OP_AND:  val[id] = add_gate(HERDB_GATE_AND, a, val[par[1]]); break;
            case HERDB_OP_OR:   val[id] = add_gate(HERDB_GATE_OR, a, val[par[1]]); break;
            case HERDB_OP_NAND: val[id] = add_gate(HERDB_GATE_NAND, a, val[par[1]]); break;
            case HERDB_OP_XOR: {   // OpenFHE v1.0.x EvalBinGate(XOR): OR(AND(a, !b), AND(!a, b))
                const Value b = val[par[1]];
                if (nd.flags & HERDB_NODE_FAST) { val[id] = add_gate(HERDB_GATE_XOR_FAST, a, b); break; }   // BINGATE XOR_FAST, one bootstrap
                Value g1 = add_gate(HERDB_GATE_AND, a, negate(b));
                Value g2 = add_gate(HERDB_GATE_AND, negate(a), b);
                val[id] = add_gate(HERDB_GATE_OR, g1, g2);
                break; }
            case HERDB_OP_MUX: {   // (sel, x, y): OR(AND(sel, x), AND(!sel, y))
                Value g1 = add_gate(HERDB_GATE_AND, a, val[par[1]]);
                Value g2 = add_gate(HERDB_GATE_AND, negate(a), val[par[2]]);
                val[id] = add_gate(HERDB_GATE_OR, g1, g2);
                break; }
            }
            break; }
        default: err = "unknown node kind"; return HERDB_ERR_ARG;
        }
    }
    for (u32 o = 0; o < out.out_bits; ++o)
        if (!out_set[o]) { err = "output bit without an output node"; return HERDB_ERR_ARG; }

    // liveness: last level reading each gate result
    for (auto& g : gates) {
        if (g.a.pool == kPoolSlots && !(g.a.flags & kOpConst)) gates[g.a.idx].last_use = std::max(gates[g.a.idx].last_use, g.level);
        if (g.b.pool == kPoolSlots && !(g.b.flags & kOpConst)) gates[g.b.idx].last_use = std::max(gates[g.b.idx].last_use, g.level);
    }
    for (auto& v : out_val)
        if (v.pool == kPoolSlots && !(v.flags & kOpConst)) gates[v.idx].last_use = UINT32_MAX;

    u32 n_levels = 0;
    for (auto& g : gates) n_levels = std::max(n_levels, g.level);
    std::vector<std::vector<u32>> by_level(n_levels + 1);
    for (u32 i = 0; i < gates.size(); ++i) by_level[gates[i].level].push_back(i);

    // slot assignment: a slot is recycled only after the level of its last reader has completed
    std::vector<u32> free_slots;
    std::vector<std::vector<u32>> expiring(n_levels + 2);
    u32 next_slot = 0;
    for (u32 L = 1; L <= n_levels; ++L) {
        for (u32 s : expiring[L - 1]) free_slots.push_back(s);
        for (u32 gi : by_level[L]) {
            BootGate& g = gates[gi];
            if (!free_slots.empty()) { g.slot = free_slots.back(); free_slots.pop_back(); }
            else g.slot = next_slot++;
            const u32 lu = g.last_use == 0 ? L : g.last_use;         // unused result: dies with its own level
            if (lu != UINT32_MAX) expiring[std::min(lu, n_levels + 1)].push_back(g.slot);
        }
    }
    out.slots = next_slot;
    out.boot_gates = (u32)gates.size();

    auto to_desc = [&](const Value& v, uint8_t& pool, u32& idx, uint8_t& flags) {
        flags = v.flags;
        if (v.flags & kOpConst) { pool = kPoolIn; idx = 0; }
        else if (v.pool == kPoolSlots) { pool = kPoolSlots; idx = gates[v.idx].slot; }
        else { pool = kPoolIn; idx = v.idx; }
    };
    out.levels.resize(n_levels);
    for (u32 L = 1; L <= n_levels; ++L) {
        for (u32 gi : by_level[L]) {
            const BootGate& g = gates[gi];
            GateDesc d{};
            d.gate = g.gate;
            to_desc(g.a, d.pool_a, d.idx_a, d.flag_a);
            to_desc(g.b, d.pool_b, d.idx_b, d.flag_b);
            d.pool_out = kPoolSlots; d.idx_out = g.slot;
            out.levels[L - 1].push_back(d);
        }
        out.max_level_width = std::max<u32>(out.max_level_width, (u32)out.levels[L - 1].size());
    }
    out.outs.resize(out.out_bits);
    for (u32 o = 0; o < out.out_bits; ++o) {
        OutDesc d{};
        to_desc(out_val[o], d.pool, d.idx, d.flags);
        out.outs[o] = d;
    }
    return HERDB_OK;
}

}  // namespace herdb
