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
constexpr int NUM_COUNTERS = 16;        // "last CTA done" tickets, one per producer kernel of a pass

struct NetWs {
    float *lut;                          // [168,16] ELU(Linear(cat[E_chrg,E_pdg,E_pv])) per combination
    float *enc;                          // [N,32]  ELU(encode_all) before bn_all
    float *emb[GM_MAX_CONV_DEPTH + 1];   // emb[0] = bn_all(enc); emb[l+1] = emb[l] + BN_l(conv_l)
    float *conv[GM_MAX_CONV_DEPTH];      // [N,32] raw EdgeConv output
    int *arg[GM_MAX_CONV_DEPTH];         // [N,32] arg-max source node per (target, channel)
    float *Y, *P;                        // [N,32] per-layer scratch: Y = W2 x, P = (W1-W2) x + b
    float *bnstat;                       // [depth+1][4][32]
    float *partials;                     // [MAX_PARTIAL_CTAS][3][32]
    int *counters;                       // [NUM_COUNTERS]
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
    t.counters = (int *)take(NUM_COUNTERS * sizeof(int));
    t.enc = (float *)take(nodes * H * sizeof(float));
    for (int l = 0; l <= depth; ++l) t.emb[l] = (float *)take(nodes * H * sizeof(float));
    for (int l = 0; l < depth; ++l) {
        t.conv[l] = (float *)take(nodes * H * sizeof(float));
        t.arg[l] = save ? (int *)take(nodes * H * sizeof(int)) : nullptr;
    }
    t.Y = (float *)take(nodes * H * sizeof(float));
    t.P = (float *)take(nodes * H * sizeof(float));
    if (w) *w = t;
    return off;
}

// split of a [32,64] EdgeConv weight row: W1 = cols 0..31 (acts on x_i), W2 = cols 32..63 (on x_j - x_i)
constexpr int CONV_IN = 2 * H;

// "Last CTA done" ticket: every CTA calls this after its global writes; exactly one CTA (the last to
// arrive) gets true and may then read what all the others wrote.  No CTA ever waits on another, so it is
// safe under any scheduling.  The counter is left at 0 for the next use.
__device__ __forceinline__ bool last_cta_done(int *counter) {
    __shared__ int s_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const int ticket = atomicAdd(counter, 1);
        const int last = ticket == (int)gridDim.x - 1;
        if (last) *counter = 0;
        s_last = last;
    }
    __syncthreads();
    const bool last = s_last != 0;
    if (last) __threadfence();
    return last;
}

// Deterministic combination of per-CTA BatchNorm moments by ONE CTA (blockDim a multiple of 32, <= 1024).
// Eight lanes cover one partial (float4 of n / mean / M2 for 4 channels each), so blockDim/8 lane groups
// stream the partials g, g+G, ... with independent 16-byte L2 loads; the G group results are then folded
// in group order by 32 threads.  The order depends only on (n_partials, blockDim): bit-deterministic.
// ATen semantics: biased variance normalises, unbiased variance feeds running_var,
// momentum update, num_batches_tracked += 1.   red: [blockDim/8][3][H] floats of shared memory.
__device__ __forceinline__ void bn_finalize_cta(const float *__restrict__ partials, int n_partials, GmBatchNorm bn,
                                                float momentum, float eps, float *__restrict__ stat,
                                                float (*red)[3][H]) {
    const int sub = threadIdx.x & 7, grp = threadIdx.x >> 3, G = blockDim.x >> 3;
    Moments m[4] = {{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}};
#pragma unroll 2
    for (int b = grp; b < n_partials; b += G) {
        const float4 *p = reinterpret_cast<const float4 *>(partials + (size_t)b * MOM_STRIDE);
        const float4 n4 = __ldcg(p + sub), me4 = __ldcg(p + 8 + sub), m24 = __ldcg(p + 16 + sub);
        m[0] = moments_combine(m[0], Moments{n4.x, me4.x, m24.x});
        m[1] = moments_combine(m[1], Moments{n4.y, me4.y, m24.y});
        m[2] = moments_combine(m[2], Moments{n4.z, me4.z, m24.z});
        m[3] = moments_combine(m[3], Moments{n4.w, me4.w, m24.w});
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        red[grp][0][4 * sub + q] = m[q].n; red[grp][1][4 * sub + q] = m[q].mean; red[grp][2][4 * sub + q] = m[q].m2;
    }
    __syncthreads();
    if (threadIdx.x < H) {
        const int c = threadIdx.x;
        Moments tot{0.f, 0.f, 0.f};
        for (int k = 0; k < G; ++k) tot = moments_combine(tot, Moments{red[k][0][c], red[k][1][c], red[k][2][c]});
        const float var = tot.n > 0.f ? tot.m2 / tot.n : 0.f;
        const float invstd = 1.0f / sqrtf(var + eps);
        stat[c] = tot.mean;
        stat[H + c] = bn.weight[c] * invstd;
        stat[2 * H + c] = invstd;
        stat[3 * H + c] = bn.bias[c];
        if (bn.running_mean) {
            const float unbiased = tot.n > 1.f ? tot.m2 / (tot.n - 1.f) : var;
            bn.running_mean[c] = (1.f - momentum) * bn.running_mean[c] + momentum * tot.mean;
            bn.running_var[c] = (1.f - momentum) * bn.running_var[c] + momentum * unbiased;
        }
        if (c == 0 && bn.num_batches_tracked) *bn.num_batches_tracked += 1;
    }
    __syncthreads();
}

// Per-layer DYNAMIC k-NN graphs (row f4: model/graph_met_network.py:63, knn_graph(emb, k, batch, loop=True) rebuilt from
// every layer's input embedding): CSR by target and, for the backward, its transpose, kept per layer.
struct KnnLayers {
    int *rowptr[GM_MAX_CONV_DEPTH], *col[GM_MAX_CONV_DEPTH], *rowptr_t[GM_MAX_CONV_DEPTH], *col_t[GM_MAX_CONV_DEPTH];
    char *knn_ws, *csr_ws;
    size_t knn_bytes, csr_bytes;
    long long ecap;
};
inline size_t knn_layers_layout(int N, int depth, int k, char *base, KnnLayers *w) {
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return base ? base + o : (char *)nullptr; };
    const size_t n = (size_t)(N > 0 ? N : 1);
    KnnLayers t{};
    t.ecap = (long long)n * k;
    for (int l = 0; l < depth; ++l) {
        t.rowptr[l] = (int *)take((n + 1) * 4);
        t.col[l] = (int *)take((size_t)t.ecap * 4);
        t.rowptr_t[l] = (int *)take((n + 1) * 4);
        t.col_t[l] = (int *)take((size_t)t.ecap * 4);
    }
    t.knn_bytes = gm_knn_graph_ws_bytes((int)n, k, 1);
    t.knn_ws = take(t.knn_bytes);
    t.csr_bytes = gm_csr_ws_bytes((int)n, t.ecap);
    t.csr_ws = take(t.csr_bytes);
    if (w) *w = t;
    return off;
}

struct AggrTail {   // BatchNorm epilogue of the EdgeConv output (train mode); partials == NULL otherwise
    float *partials; int *counter; GmBatchNorm bn; float momentum, eps; float *stat;
};

// ----------------------------------------------------------------------------------------- BatchNorm-backward plumbing
// What a consumer needs to turn G into d(raw) for one BatchNorm: per channel
//     d(raw) = scale * (G - c1 - (raw - mean) * invstd * c2),  c1 = sum(G)/N, c2 = sum(G*xhat)/N  (train)
//                                                              c1 = c2 = 0                        (eval)
struct BnBwd {
    const float *stat;      // mean[32], scale[32], invstd[32], beta[32]   (forward)
    const float *sum;       // sum(G)[32], sum(G*xhat)[32]                 (from the producer's last CTA)
    float inv_n;            // 1/N in train mode, 0 in eval mode
};
// Producer side: where the per-CTA partial sums go and what the last CTA writes.
struct BnSumOut {
    const float *raw;       // [N,32] pre-BatchNorm activations of the BatchNorm being differentiated
    const float *stat;
    float *part;            // [grid][64]
    float *sum;             // [64]
    float *dgamma, *dbeta;  // gradient tensors (may be NULL)
    int *counter;
};

// smem coefficient table for BnBwd: [5][32] = mean, invstd, scale, c1, c2
__device__ __forceinline__ void bn_bwd_coeffs(const BnBwd &b, float *s_coef) {
    const int t = threadIdx.x;
    if (t < H) {
        s_coef[t] = b.stat[t];
        s_coef[H + t] = b.stat[2 * H + t];
        s_coef[2 * H + t] = b.stat[H + t];
        s_coef[3 * H + t] = b.sum[t] * b.inv_n;
        s_coef[4 * H + t] = b.sum[H + t] * b.inv_n;
    }
}
__device__ __forceinline__ float bn_bwd_one(const float *s_coef, int c, float g, float raw) {
    const float xhat = (raw - s_coef[c]) * s_coef[H + c];
    return s_coef[2 * H + c] * (g - s_coef[3 * H + c] - xhat * s_coef[4 * H + c]);
}


}  // namespace gm
