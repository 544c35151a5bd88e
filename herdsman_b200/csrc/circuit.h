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

struct CompiledCircuit {
    u32 in_bits = 0, out_bits = 0;
    u32 boot_gates = 0;
    u32 slots = 0;                                  // ciphertext slots per row (after liveness reuse)
    u32 max_level_width = 0;
    std::vector<std::vector<GateDesc>> levels;      // levels[l] = independent gates
    std::vector<OutDesc> outs;                      // one per output bit
};

// returns HERDB_OK or HERDB_ERR_ARG (err filled)
int compile_circuit(const herdb_circuit* c, CompiledCircuit& out, std::string& err);

}  // namespace herdb
