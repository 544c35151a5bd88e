// crosscheck.cpp -- what would PIN parity: OpenFHE generates the keys and evaluates the gates, the keys
// are exported into this repo's flat blob (include/herdb_params.h), the oracle evaluates the same gates on
// the same ciphertexts, and every output ciphertext must be identical.
//
// STATUS: written against the OpenFHE v1.0.3 binfhe headers from memory; it has NOT been compiled -- OpenFHE
// is not installed in the build image and cannot be fetched (no network).  It documents the mapping between
// OpenFHE's objects and the flat formats, which is the part that matters for an importer:
//
//   RingGSWACCKey ek = cc.GetRefreshKey()        (GINX: dimensions [1][2][n])
//     (*ek)[0][s][i]         RGSW(s_i == +1 ? 1 : 0) for s = 0, RGSW(s_i == -1 ? 1 : 0) for s = 1
//     (*key)[row][col]       NativePoly in EVALUATION format, row = 0 .. 2*digitsG-1, col = 0 (a), 1 (b)
//       -> blob BSK u32 [i][s][row][col][N] after SetFormat(COEFFICIENT)
//   LWESwitchingKey ksk = cc.GetSwitchKey()
//     GetElementsA()[i][v][j]  NativeVector(n),  GetElementsB()[i][v][j]  NativeInteger     i < N, v < baseKS, j < digitsKS
//       -> blob KSK u16 [i][v][j][n+1]  (a[0..n) then b)
//   LWECiphertext ct: GetA() NativeVector(n), GetB() NativeInteger, modulus q
//       -> (n+1) u16
#include <cstdio>
#include <cstring>
#include <vector>

#include "binfhecontext.h"

extern "C" {
#include "../../oracle/oracle.h"
}
#include "herdb_params.h"

using namespace lbcrypto;

static std::vector<uint16_t> flat(ConstLWECiphertext ct) {
    const auto& a = ct->GetA();
    std::vector<uint16_t> out(a.GetLength() + 1);
    for (size_t k = 0; k < a.GetLength(); ++k) out[k] = (uint16_t)a[k].ConvertToInt();
    out[a.GetLength()] = (uint16_t)ct->GetB().ConvertToInt();
    return out;
}

int main() {
    auto cc = BinFHEContext();
    cc.GenerateBinFHEContext(STD128, GINX);
    auto sk = cc.KeyGen();
    cc.BTKeyGen(sk);
    const auto params = cc.GetParams();
    const uint32_t n = params->GetLWEParams()->Getn(), N = params->GetLWEParams()->GetN();
    const uint32_t dG = params->GetRingGSWParams()->GetDigitsG();
    const uint32_t baseKS = params->GetLWEParams()->GetBaseKS();
    const uint32_t dKS = 2;   // ceil(log(qKS) / log(baseKS)) for STD128
    if (n != HERDB_N_STD128 || N != HERDB_RING_N || dG != HERDB_DIGITS_G ||
        params->GetLWEParams()->GetQ().ConvertToInt() != HERDB_Q || params->GetLWEParams()->Getq().ConvertToInt() != HERDB_LWE_Q) {
        std::printf("parameter set differs from include/herdb_params.h -- correct the header first\n");
        return 2;
    }
    // ---- export the keys into the flat blob
    std::vector<uint8_t> blob(sizeof(herdb_key_header) + (size_t)n * 2 * 2 * dG * 2 * N * 4 + (size_t)N * baseKS * dKS * (n + 1) * 2);
    herdb_key_header h{};
    h.magic = HERDB_KEY_MAGIC; h.version = HERDB_KEY_VERSION; h.n = n; h.ring_n = N; h.q = HERDB_LWE_Q; h.Q = HERDB_Q; h.qKS = HERDB_QKS;
    h.baseG_log = HERDB_BASEG_LOG; h.digitsG = dG; h.baseKS_log = HERDB_BASEKS_LOG; h.digitsKS = dKS;
    std::memcpy(blob.data(), &h, sizeof h);
    uint32_t* bsk = reinterpret_cast<uint32_t*>(blob.data() + sizeof h);
    const auto ek = cc.GetRefreshKey();
    for (uint32_t i = 0; i < n; ++i)
        for (uint32_t s = 0; s < 2; ++s) {
            const auto& key = (*ek)[0][s][i];
            for (uint32_t row = 0; row < 2 * dG; ++row)
                for (uint32_t col = 0; col < 2; ++col) {
                    NativePoly p = (*key)[row][col];
                    p.SetFormat(Format::COEFFICIENT);
                    uint32_t* dst = bsk + ((((size_t)i * 2 + s) * 2 * dG + row) * 2 + col) * N;
                    for (uint32_t k = 0; k < N; ++k) dst[k] = (uint32_t)p[k].ConvertToInt();
                }
        }
    uint16_t* ksk = reinterpret_cast<uint16_t*>(bsk + (size_t)n * 2 * 2 * dG * 2 * N);
    const auto sw = cc.GetSwitchKey();
    for (uint32_t i = 0; i < N; ++i)
        for (uint32_t v = 0; v < baseKS; ++v)
            for (uint32_t j = 0; j < dKS; ++j) {
                uint16_t* dst = ksk + (((size_t)i * baseKS + v) * dKS + j) * (n + 1);
                const auto& a = sw->GetElementsA()[i][v][j];
                for (uint32_t k = 0; k < n; ++k) dst[k] = (uint16_t)a[k].ConvertToInt();
                dst[n] = (uint16_t)sw->GetElementsB()[i][v][j].ConvertToInt();
            }
    orc_ctx* orc = orc_create();
    if (orc_load_key(orc, blob.data(), blob.size()) != 0) { std::printf("oracle rejected the exported key\n"); return 2; }
    // ---- same gates on the same ciphertexts
    const BINGATE gates[] = {OR, AND, NOR, NAND, XOR_FAST, XNOR_FAST};
    size_t bad = 0, total = 0;
    for (int g = 0; g < 6; ++g)
        for (int m1 = 0; m1 < 2; ++m1)
            for (int m2 = 0; m2 < 2; ++m2) {
                auto c1 = cc.Encrypt(sk, m1), c2 = cc.Encrypt(sk, m2);
                auto ref = flat(cc.EvalBinGate(gates[g], c1, c2));
                auto a = flat(c1), b = flat(c2);
                std::vector<uint16_t> got(n + 1);
                orc_eval_gate_batch(orc, g, a.data(), b.data(), got.data(), 1, 1);
                ++total;
                if (got != ref) { ++bad; std::printf("gate %d inputs (%d, %d): oracle != OpenFHE\n", g, m1, m2); }
            }
    std::printf("%zu / %zu gate outputs identical to OpenFHE -- parity %s\n", total - bad, total, bad ? "NOT pinned" : "PINNED");
    return bad ? 1 : 0;
}
