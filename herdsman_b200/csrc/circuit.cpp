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

inline bool gate_operand(const Value& v) { return v.pool == kPoolSlots && !(v.flags & kOpConst); }

// ---- level scheduling ------------------------------------------------------------------------------------
// Every gate may sit on any level between its as-soon-as-possible and its as-late-as-possible one (critical
// path length fixed).  A level of g gates costs ceil(rows * g / wave) rounds of the blind-rotation grid, so the
// last round of a level is usually part empty; gates with slack are moved to fill those rounds.  Several
// deadline-driven list schedules are built and the cheapest kept; ties keep the as-soon-as-possible levels,
// so a hint can only lower the cost.  The result is the same ciphertexts: a gate's output does not depend on
// when it runs.
uint64_t rounds_of(uint64_t gates, const SchedHint& h) { return (gates * h.rows + h.wave - 1) / h.wave; }

uint64_t schedule_cost(const std::vector<u32>& level, u32 depth, const SchedHint& h) {
    std::vector<u32> width(depth + 1, 0);
    for (u32 l : level) width[l]++;
    uint64_t c = 0;
    for (u32 L = 1; L <= depth; ++L) c += rounds_of(width[L], h);
    return c;
}

// mode 1: every level takes the gates that are due, then fills the rounds those need with the most urgent
//         ready gates;  mode 2: as 1, and a level on which nothing is due still runs one full round if it can
//         be filled;  mode 3: every level takes the number of ready gates with the best fill (at least the due ones)
std::vector<u32> list_schedule(const std::vector<BootGate>& gates, const std::vector<std::vector<u32>>& children,
                               const std::vector<u32>& alap, u32 depth, const SchedHint& h, int mode) {
    const u32 ng = (u32)gates.size();
    std::vector<u32> level(ng, 0), pending(ng, 0);
    std::vector<std::vector<u32>> bucket(depth + 2);
    for (u32 i = 0; i < ng; ++i) {
        pending[i] = (gate_operand(gates[i].a) ? 1u : 0u) + (gate_operand(gates[i].b) ? 1u : 0u);
        if (!pending[i]) bucket[1].push_back(i);
    }
    typedef std::pair<u32, u32> Item;                                   // (deadline, gate)
    std::priority_queue<Item, std::vector<Item>, std::greater<Item>> ready;
    std::vector<u32> take;
    for (u32 L = 1; L <= depth; ++L) {
        for (u32 g : bucket[L]) ready.push(Item(alap[g], g));
        take.clear();
        while (!ready.empty() && ready.top().first <= L) { take.push_back(ready.top().second); ready.pop(); }
        const uint64_t due = take.size(), avail = due + ready.size();
        uint64_t want = due;
        if (mode == 3) {
            double best = -1.0;
            for (uint64_t k = std::max<uint64_t>(due, 1); k <= avail; ++k) {
                const double fill = (double)(k * h.rows) / (double)(rounds_of(k, h) * h.wave);
                if (fill >= best) { best = fill; want = k; }
            }
            if (!due && best < 0.97) want = 0;
        } else {
            if (due) want = rounds_of(due, h) * h.wave / h.rows;
            else if (mode == 2 && avail * h.rows >= h.wave) want = h.wave / h.rows;
        }
        while (take.size() < want && !ready.empty()) { take.push_back(ready.top().second); ready.pop(); }
        for (u32 g : take) {
            level[g] = L;
            for (u32 ch : children[g]) if (--pending[ch] == 0) bucket[L + 1].push_back(ch);
        }
    }
    return level;
}

// Local search on a schedule: move single gates inside their dependency window while that strictly lowers the
// cost, then try to empty the last round of a level by moving exactly the gates that spill into it to levels
// with room in their last round.
void improve_schedule(std::vector<u32>& level, const std::vector<BootGate>& gates, const std::vector<std::vector<u32>>& children,
                      u32 depth, const SchedHint& h) {
    const u32 ng = (u32)gates.size();
    std::vector<u32> width(depth + 2, 0);
    for (u32 l : level) width[l]++;
    auto window = [&](u32 g, u32& lo, u32& hi) {
        lo = 1; hi = depth;
        if (gate_operand(gates[g].a)) lo = std::max(lo, level[gates[g].a.idx] + 1);
        if (gate_operand(gates[g].b)) lo = std::max(lo, level[gates[g].b.idx] + 1);
        for (u32 ch : children[g]) hi = std::min(hi, level[ch] - 1);
    };
    auto delta = [&](u32 from, u32 to) -> long long {
        return (long long)rounds_of(width[from] - 1, h) - (long long)rounds_of(width[from], h) +
               (long long)rounds_of(width[to] + 1, h) - (long long)rounds_of(width[to], h);
    };
    for (int pass = 0; pass < 64; ++pass) {
        bool moved = false;
        for (u32 g = 0; g < ng; ++g) {                                  // single moves
            u32 lo, hi; window(g, lo, hi);
            const u32 A = level[g];
            u32 bestB = A; long long bestD = 0;
            for (u32 B = lo; B <= hi; ++B) {
                if (B == A) continue;
                const long long d = delta(A, B);
                if (d < bestD) { bestD = d; bestB = B; }
            }
            if (bestB != A) { width[A]--; width[bestB]++; level[g] = bestB; moved = true; }
        }
        std::vector<std::vector<u32>> members(depth + 1);
        for (u32 g = 0; g < ng; ++g) members[level[g]].push_back(g);
        for (u32 A = 1; A <= depth; ++A) {                              // shed the spill of level A
            if (!width[A]) continue;
            const uint64_t keep = (rounds_of(width[A], h) - 1) * h.wave / h.rows;
            const u32 need = width[A] - (u32)std::min<uint64_t>(keep, width[A]);
            if (!need || need > 64) continue;
            std::vector<std::pair<u32, u32>> done;                      // (gate, old level)
            for (u32 g : members[A]) {
                if (level[g] != A) continue;
                u32 lo, hi; window(g, lo, hi);
                for (u32 B = lo; B <= hi; ++B) {
                    if (B == A || !width[B]) continue;
                    if (rounds_of(width[B] + 1, h) != rounds_of(width[B], h)) continue;
                    done.push_back(std::make_pair(g, A));
                    width[A]--; width[B]++; level[g] = B;
                    break;
                }
                if (done.size() == need) break;
            }
            if (done.size() == need) moved = true;
            else for (auto& m : done) { width[level[m.first]]--; width[m.second]++; level[m.first] = m.second; }
        }
        if (!moved) break;
    }
}

// gates[].level: as-soon-as-possible on entry, the chosen schedule on return (levels renumbered without gaps)
void schedule_levels(std::vector<BootGate>& gates, const std::vector<Value>& out_val, const SchedHint& h, CompiledCircuit& out) {
    const u32 ng = (u32)gates.size();
    if (!h.rows || !h.wave || !ng) return;
    u32 depth = 0;
    for (auto& g : gates) depth = std::max(depth, g.level);
    std::vector<std::vector<u32>> children(ng);
    for (u32 i = 0; i < ng; ++i) {
        if (gate_operand(gates[i].a)) children[gates[i].a.idx].push_back(i);
        if (gate_operand(gates[i].b)) children[gates[i].b.idx].push_back(i);
    }
    (void)out_val;
    std::vector<u32> asap(ng), alap(ng, depth);
    for (u32 i = 0; i < ng; ++i) asap[i] = gates[i].level;
    for (u32 i = ng; i-- > 0;)                                           // gates are in topological order
        for (u32 ch : children[i]) alap[i] = std::min(alap[i], alap[ch] - 1);
    std::vector<u32> best = asap;
    out.rounds_asap = out.rounds = schedule_cost(asap, depth, h);
    auto offer = [&](std::vector<u32> cand, u32 policy) {
        improve_schedule(cand, gates, children, depth, h);
        const uint64_t c = schedule_cost(cand, depth, h);
        if (c < out.rounds) { out.rounds = c; best = cand; out.policy = policy; }
    };
    for (int mode = 1; mode <= 3; ++mode) offer(list_schedule(gates, children, alap, depth, h, mode), (u32)mode);
    offer(alap, 4);
    offer(asap, 5);
    std::vector<u32> used(depth + 2, 0), renum(depth + 2, 0);
    for (u32 l : best) used[l] = 1;
    u32 next = 0;
    for (u32 L = 1; L <= depth; ++L) if (used[L]) renum[L] = ++next;
    for (u32 i = 0; i < ng; ++i) gates[i].level = renum[best[i]];
}

Value negate(Value v) {
    if (v.flags & kOpConst) v.flags ^= kOpConstVal; else v.flags ^= kOpNeg;
    return v;
}

}  // namespace

int compile_circuit(const herdb_circuit* c, CompiledCircuit& out, std::string& err, SchedHint hint) {
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

    schedule_levels(gates, out_val, hint, out);

    // liveness: last level reading each gate result
    for (auto& g : gates) {
        if (gate_operand(g.a)) gates[g.a.idx].last_use = std::max(gates[g.a.idx].last_use, g.level);
        if (gate_operand(g.b)) gates[g.b.idx].last_use = std::max(gates[g.b.idx].last_use, g.level);
    }
    for (auto& v : out_val)
        if (gate_operand(v)) gates[v.idx].last_use = UINT32_MAX;

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

int eval_compiled_plain(const CompiledCircuit& cc, const uint8_t* in_bits, uint8_t* out_bits) {
    std::vector<uint8_t> slots(cc.slots, 0), fresh;
    std::vector<u32> read_at(cc.slots, 0), written_at(cc.slots, 0);       // level (1-based) of the last read / write
    auto operand = [&](uint8_t pool, u32 idx, uint8_t flags) -> uint8_t {
        if (flags & kOpConst) return (flags & kOpConstVal) ? 1 : 0;
        const uint8_t v = pool == kPoolSlots ? slots[idx] : in_bits[idx];
        return (flags & kOpNeg) ? (uint8_t)(v ^ 1) : v;
    };
    u32 lev = 0;
    for (auto& L : cc.levels) {
        ++lev;
        fresh.clear();
        for (auto& g : L) {                                  // a level reads the slot file as the previous level left it
            if (!(g.flag_a & kOpConst) && g.pool_a == kPoolSlots) read_at[g.idx_a] = lev;
            if (!(g.flag_b & kOpConst) && g.pool_b == kPoolSlots) read_at[g.idx_b] = lev;
            const uint8_t a = operand(g.pool_a, g.idx_a, g.flag_a), b = operand(g.pool_b, g.idx_b, g.flag_b);
            uint8_t r = 0;
            switch (g.gate) {
            case HERDB_GATE_OR: r = a | b; break;
            case HERDB_GATE_AND: r = a & b; break;
            case HERDB_GATE_NOR: r = (a | b) ^ 1; break;
            case HERDB_GATE_NAND: r = (a & b) ^ 1; break;
            case HERDB_GATE_XOR_FAST: r = a ^ b; break;
            case HERDB_GATE_XNOR_FAST: r = a ^ b ^ 1; break;
            }
            fresh.push_back(r);
        }
        for (size_t k = 0; k < L.size(); ++k) {              // gates of a level run concurrently on the device:
            const u32 o = L[k].idx_out;                      // none may write a slot the level reads or writes
            if (read_at[o] == lev || written_at[o] == lev) return -1;
            written_at[o] = lev;
            slots[o] = fresh[k];
        }
    }
    for (u32 o = 0; o < cc.out_bits; ++o) out_bits[o] = operand(cc.outs[o].pool, cc.outs[o].idx, cc.outs[o].flags);
    return 0;
}

}  // namespace herdb
