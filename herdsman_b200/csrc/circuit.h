// circuit.h -- host-side circuit compiler: herd::common::Circuit (as the flat herdb_circuit of
// include/herdb.h) -> levels of bootstrapped gates over a per-row ciphertext slot file.
//
// What the reference worker does per row, node by node, is re-ordered here
// into "all gates of one dependency level for all rows in one launch".  NOT and constants cost
// no bootstrap (OpenFHE EvalNOT / trivial ciphertext): they are folded into operand flags.
// XOR and MUX (operation.hpp:12-20; not expressible on the wire, node.proto:25-30) are lowered
// to the AND/OR/NOT form OpenFHE v1.0.x itself uses for EvalBinGate(XOR).
#pragma once
#include <string>
#include <vector>
#include "../../include/herdb.h"
#include "br_core.cuh"
#include "kernels.h"

namespace herdb {

enum PoolId : uint8_t { kPoolIn = 0, kPoolSlots = 1, kPoolAux = 2, kPoolOut = 3 };

// What one launch looks like to the level scheduler: `rows` rows go through every level together, and the
// blind-rotation grid retires `wave` jobs (gates x rows) per round of its persistent CTAs (SM count x
// ciphertexts per CTA).  A level of g gates therefore costs ceil(rows * g / wave) rounds, and gates with slack
// are worth moving to the level whose last round they fill.  rows == 0: keep the as-soon-as-possible levels.
struct SchedHint { uint64_t rows = 0; u32 wave = 0; };
constexpr u32 kBrCtsPerCta = 4;                     // blind_rotate_kernel<4> (kernels.cu)

struct CompiledCircuit {
    u32 in_bits = 0, out_bits = 0;
    u32 boot_gates = 0;
    u32 slots = 0;                                  // ciphertext slots per row (after liveness reuse)
    u32 max_level_width = 0;
    std::vector<std::vector<GateDesc>> levels;      // levels[l] = independent gates
    std::vector<OutDesc> outs;                      // one per output bit
    uint64_t rounds_asap = 0, rounds = 0;           // cost of the as-soon-as-possible levels / of the chosen ones (0 without a hint)
    u32 policy = 0;                                 // which schedule won (0 = as soon as possible)
};

// returns HERDB_OK or HERDB_ERR_ARG (err filled)
int compile_circuit(const herdb_circuit* c, CompiledCircuit& out, std::string& err, SchedHint hint = SchedHint());

// plaintext interpreter of a compiled circuit (levels in order, slot file with reuse): what the device computes,
// on bits.  Host-side check of the compiler; not on any product path.  -1 if two gates of one level touch the
// same slot in conflicting ways (a race on the device), else 0.
int eval_compiled_plain(const CompiledCircuit& cc, const uint8_t* in_bits, uint8_t* out_bits);

}  // namespace herdb
