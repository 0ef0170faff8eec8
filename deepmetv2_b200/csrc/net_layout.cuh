// Workspace layout shared by gm_net_forward / gm_net_backward.
#pragma once
#include "common.cuh"

namespace gm {

constexpr int NODE_TILE = 128;          // nodes per CTA tile in the node-level kernels
constexpr int TILE_LD = H + 1;          // padded smem row (33 floats): conflict-free row-per-thread access
constexpr int LUT_ROWS = 3 * 7 * 8;     // (charge, pdg, fromPV) combinations
constexpr int MAX_PARTIAL_CTAS = 148 * 8;
constexpr int MOM_STRIDE = 3 * H;       // per-CTA BatchNorm partial: n[32], mean[32], m2[32]
constexpr int BNSTAT_STRIDE = 4 * H;    // per BN layer: mean[32], scale(=gamma*invstd)[32], invstd[32], beta[32]

struct NetWs {
    float *lut;                          // [168,16] ELU(Linear(cat[E_chrg,E_pdg,E_pv])) per combination
    float *enc;                          // [N,32]  ELU(encode_all) before bn_all
    float *emb[GM_MAX_CONV_DEPTH + 1];   // emb[0] = bn_all(enc); emb[l+1] = emb[l] + BN_l(conv_l)
    float *conv[GM_MAX_CONV_DEPTH];      // [N,32] raw EdgeConv output
    int *arg[GM_MAX_CONV_DEPTH];         // [N,32] arg-max source node per (target, channel)
    float *bnstat;                       // [depth+1][4][32]
    float *partials;                     // [MAX_PARTIAL_CTAS][3][32]
};

inline size_t net_ws_layout(int N, int depth, int save, char *base, NetWs *w) {
    size_t off = 0;
    auto take = [&](size_t bytes) {
        size_t o = off;
        off = align_up(off + bytes, 256);
        return base ? base + o : nullptr;
    };
    const size_t nodes = (size_t)(N > 0 ? N : 1);
    NetWs t{};
    t.lut = (float *)take(LUT_ROWS * 16 * sizeof(float));
    t.bnstat = (float *)take((size_t)(GM_MAX_CONV_DEPTH + 1) * BNSTAT_STRIDE * sizeof(float));
    t.partials = (float *)take((size_t)MAX_PARTIAL_CTAS * MOM_STRIDE * sizeof(float));
    t.enc = (float *)take(nodes * H * sizeof(float));
    for (int l = 0; l <= depth; ++l) t.emb[l] = (float *)take(nodes * H * sizeof(float));
    for (int l = 0; l < depth; ++l) {
        t.conv[l] = (float *)take(nodes * H * sizeof(float));
        t.arg[l] = save ? (int *)take(nodes * H * sizeof(int)) : nullptr;
    }
    if (w) *w = t;
    return off;
}

// split of a [32,64] EdgeConv weight row: W1 = cols 0..31 (acts on x_i), W2 = cols 32..63 (on x_j - x_i)
constexpr int CONV_IN = 2 * H;

}  // namespace gm
