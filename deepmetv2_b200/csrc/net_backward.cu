// Backward of the GraphMETNetwork hot path (what loss.backward() runs through autograd in the
// reference, train.py:51): hand-written, atomics-free, fixed-order (bit-deterministic) kernels.
//
//   k_head_bwd        : d(out) -> G = d(emb_L), partial dW/db of output.{0,2}, BatchNorm sums of the last layer
//   k_scatter_s       : S[j][c] = sum_{i: argmax_src(i,c) == j} d(conv)[i][c]  as a GATHER over the
//                       transposed CSR (max-aggregation routes each (target,channel) gradient to one edge);
//                       d(conv) = BatchNorm backward of G, evaluated on the fly for the matching entries only
//   k_conv_bwd_node2  : G += (W1-W2)^T d(conv) + W2^T S ; partial dW1 = dconv (x) x, dW2 = (S-dconv) (x) x ;
//                       BatchNorm sums of the layer below
//   k_encode_bwd      : BN_all backward on the fly -> encode_all / embed_continuous / embed_categorical / 3 tables
//   k_reduce_partials : ONE launch: fixed-order sum of every stage's per-CTA partials into the gradient tensors
//
// BatchNorm backward needs sum(G) and sum(G*xhat) over all nodes before any node can be finished.  The
// kernel that PRODUCES G accumulates per-CTA partial sums and its last CTA (ticket counter, nobody waits)
// folds them in a fixed order into bnsum (and d gamma / d beta), so the consumer kernels apply
//     d(conv) = gamma*invstd * (G - mean(G) - xhat * mean(G*xhat))
// while they load G: no separate BatchNorm-backward pass, no extra launch.
//
// Parameter-gradient GEMMs (dW = sum_n a_n (x) b_n) are split-K over node tiles: every CTA keeps 4x4
// register blocks of the outer product for its tiles, the per-CTA results are summed in CTA order.
#include "geo_layout.cuh"
#include "net_layout.cuh"

namespace gm {
namespace {

constexpr int GLD = 36;                 // float4-aligned padded tile row (conflict-free LDS.128 row-per-thread)
constexpr int GP_STRIDE = 2112;         // floats per CTA in the parameter-gradient partial buffer
// resident CTAs per SM the backward node kernels are compiled for (= their grid: one wave)
#ifndef GM_BWD_OCC_HEAD
#define GM_BWD_OCC_HEAD 4
#endif
#ifndef GM_BWD_OCC_ENC
#define GM_BWD_OCC_ENC 2
#endif
#ifndef GM_BWD_OCC_CONV
#define GM_BWD_OCC_CONV 3
#endif
constexpr int BWD_OCC_MAX = GM_BWD_OCC_HEAD > GM_BWD_OCC_CONV ? (GM_BWD_OCC_HEAD > GM_BWD_OCC_ENC ? GM_BWD_OCC_HEAD : GM_BWD_OCC_ENC)
                                                             : (GM_BWD_OCC_CONV > GM_BWD_OCC_ENC ? GM_BWD_OCC_CONV : GM_BWD_OCC_ENC);
constexpr int MAX_BWD_CTAS = 148 * BWD_OCC_MAX;
constexpr int MAX_STAGES = GM_MAX_CONV_DEPTH + 2;   // head, conv layers, encoder

struct BwdWs {
    float *G, *S;                       // [N,32] each
    float *bnsum;                       // [depth+1][64]  sum(G)[32], sum(G*xhat)[32] per BatchNorm
    float *bnpart;                      // [MAX_BWD_CTAS][64]
    float *gpart;                       // [MAX_STAGES][MAX_BWD_CTAS][GP_STRIDE]
    int *counters;                      // [NUM_COUNTERS]
};

size_t bwd_ws_layout(int N, char *base, BwdWs *w) {
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return base ? base + o : nullptr; };
    const size_t nodes = (size_t)(N > 0 ? N : 1);
    BwdWs t;
    t.counters = (int *)take(NUM_COUNTERS * sizeof(int));
    t.bnsum = (float *)take((size_t)(GM_MAX_CONV_DEPTH + 1) * 64 * sizeof(float));
    t.bnpart = (float *)take((size_t)MAX_BWD_CTAS * 64 * sizeof(float));
    t.gpart = (float *)take((size_t)MAX_STAGES * MAX_BWD_CTAS * GP_STRIDE * sizeof(float));
    t.G = (float *)take(nodes * H * sizeof(float));
    t.S = (float *)take(nodes * H * sizeof(float));
    if (w) *w = t;
    return off;
}

// ----------------------------------------------------------------------------------------- helpers
__device__ __forceinline__ void load_row32(const float *__restrict__ src, float (&v)[H]) {
    const float4 *s4 = reinterpret_cast<const float4 *>(src);
#pragma unroll
    for (int k = 0; k < H / 4; ++k) {
        float4 q = s4[k];
        v[4 * k] = q.x; v[4 * k + 1] = q.y; v[4 * k + 2] = q.z; v[4 * k + 3] = q.w;
    }
}
template <int NV>
__device__ __forceinline__ void store_row(float *dst, const float (&v)[NV]) {
    float4 *d4 = reinterpret_cast<float4 *>(dst);
#pragma unroll
    for (int k = 0; k < NV / 4; ++k) d4[k] = make_float4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
}

// dW[o][i] += sum_rows A[r][o] * B[r][i] ; thread owns one 4x4 block for one group of rows.
template <int O, int I, int NT = NODE_TILE>
struct Outer {
    static constexpr int BO = O / 4, BI = I / 4, BLK = BO * BI;
    static constexpr int GROUPS = NT / BLK;
    static constexpr int ROWS_PER = (NODE_TILE + GROUPS - 1) / GROUPS;
    static_assert(O % 4 == 0 && I % 4 == 0 && BLK <= NT, "bad outer-product shape");
    float acc[16];
    __device__ __forceinline__ void zero() {
#pragma unroll
        for (int k = 0; k < 16; ++k) acc[k] = 0.f;
    }
    __device__ __forceinline__ void accumulate(const float *A, int lda, const float *B, int ldb, int rows) {
        const int t = threadIdx.x;
        if (t >= GROUPS * BLK) return;
        const int g = t / BLK, b = t - g * BLK, bo = b / BI, bi = b - bo * BI;
        const int r0 = g * ROWS_PER, r1 = min(rows, r0 + ROWS_PER);
        for (int r = r0; r < r1; ++r) {
            const float4 a = *reinterpret_cast<const float4 *>(A + r * lda + 4 * bo);
            const float4 v = *reinterpret_cast<const float4 *>(B + r * ldb + 4 * bi);
            acc[0] = fmaf(a.x, v.x, acc[0]);   acc[1] = fmaf(a.x, v.y, acc[1]);
            acc[2] = fmaf(a.x, v.z, acc[2]);   acc[3] = fmaf(a.x, v.w, acc[3]);
            acc[4] = fmaf(a.y, v.x, acc[4]);   acc[5] = fmaf(a.y, v.y, acc[5]);
            acc[6] = fmaf(a.y, v.z, acc[6]);   acc[7] = fmaf(a.y, v.w, acc[7]);
            acc[8] = fmaf(a.z, v.x, acc[8]);   acc[9] = fmaf(a.z, v.y, acc[9]);
            acc[10] = fmaf(a.z, v.z, acc[10]); acc[11] = fmaf(a.z, v.w, acc[11]);
            acc[12] = fmaf(a.w, v.x, acc[12]); acc[13] = fmaf(a.w, v.y, acc[13]);
            acc[14] = fmaf(a.w, v.z, acc[14]); acc[15] = fmaf(a.w, v.w, acc[15]);
        }
    }
    // one-barrier form of flush: every row group parks its block in stage[g][O][I] (GROUPS*O*I floats); after a barrier
    // fold_staged sums the groups in the same order as flush (bit-identical)
    __device__ __forceinline__ void stage(float *stage) const {
        const int t = threadIdx.x;
        if (t >= GROUPS * BLK) return;
        const int g = t / BLK, b = t - g * BLK, bo = b / BI, bi = b - bo * BI;
#pragma unroll
        for (int a = 0; a < 4; ++a)
            *reinterpret_cast<float4 *>(stage + g * (O * I) + (4 * bo + a) * I + 4 * bi) =
                make_float4(acc[4 * a], acc[4 * a + 1], acc[4 * a + 2], acc[4 * a + 3]);
    }
    static __device__ __forceinline__ void fold_staged(const float *stage, float *__restrict__ dst) {
        for (int i = threadIdx.x; i < O * I; i += NT) {
            float s = 0.f;
#pragma unroll
            for (int g = 0; g < GROUPS; ++g) s += stage[g * (O * I) + i];
            dst[i] = s;
        }
    }
    // fold the row groups in order, write the [O][I] CTA partial; scratch >= O*I floats of smem
    __device__ __forceinline__ void flush(float *scratch, float *__restrict__ dst) {
        const int t = threadIdx.x;
        for (int i = t; i < O * I; i += NT) scratch[i] = 0.f;
        __syncthreads();
        const int g = t / BLK, b = t - g * BLK, bo = b / BI, bi = b - bo * BI;
        for (int gg = 0; gg < GROUPS; ++gg) {
            if (t < GROUPS * BLK && g == gg) {
#pragma unroll
                for (int a = 0; a < 4; ++a)
#pragma unroll
                    for (int c = 0; c < 4; ++c) scratch[(4 * bo + a) * I + 4 * bi + c] += acc[4 * a + c];
            }
            __syncthreads();
        }
        for (int i = t; i < O * I; i += NT) dst[i] = scratch[i];
        __syncthreads();
    }
};

// column sums of a tile: s[c] += sum_rows T[r][c], c < NC
template <int NC, int NT = NODE_TILE>
struct ColSum {
    static constexpr int GROUPS = NT / NC;
    static constexpr int ROWS_PER = (NODE_TILE + GROUPS - 1) / GROUPS;
    float acc;
    __device__ __forceinline__ void zero() { acc = 0.f; }
    __device__ __forceinline__ void accumulate(const float *T, int ld, int rows) {
        const int t = threadIdx.x;
        if (t >= GROUPS * NC) return;
        const int g = t / NC, c = t - g * NC;
        const int r0 = g * ROWS_PER, r1 = min(rows, r0 + ROWS_PER);
        for (int r = r0; r < r1; ++r) acc += T[r * ld + c];
    }
    __device__ __forceinline__ void stage(float *stage) const {      // stage[g][NC]; see Outer::stage
        if (threadIdx.x < GROUPS * NC) stage[threadIdx.x] = acc;
    }
    static __device__ __forceinline__ void fold_staged(const float *stage, float *__restrict__ dst, int ncols_out) {
        if ((int)threadIdx.x < ncols_out) {
            float s = 0.f;
#pragma unroll
            for (int g = 0; g < GROUPS; ++g) s += stage[g * NC + threadIdx.x];
            dst[threadIdx.x] = s;
        }
    }
    __device__ __forceinline__ void flush(float *scratch, float *__restrict__ dst, int ncols_out) {
        const int t = threadIdx.x;
        if (t < NC) scratch[t] = 0.f;
        __syncthreads();
        const int g = t / NC, c = t - g * NC;
        for (int gg = 0; gg < GROUPS; ++gg) {
            if (t < GROUPS * NC && g == gg) scratch[c] += acc;
            __syncthreads();
        }
        if (t < ncols_out) dst[t] = scratch[t];
        __syncthreads();
    }
};

struct Seg { const float *part; float *dst; int dst_ld, off, rows, src_ld, cols, nblk; };
constexpr int MAX_SEGS = 4 + 3 * GM_MAX_CONV_DEPTH + 9;
struct Segs { int n; Seg s[MAX_SEGS]; };

// dst[r*dst_ld + c] = sum_b part[b][off + r*src_ld + c].  grid = (chunks of 32 outputs, segments);
// thread (c = t%32, g = t/32): lanes read 32 consecutive outputs (coalesced), the RP_GROUPS groups split the
// CTA-partial axis (b = g, g+RP_GROUPS, ...) and are folded in a fixed order => deterministic.  32 groups: the launch is
// bound by the latency of each thread's dependent batches of loads (444 partials / 8 groups = 14 batches of 4 before).
constexpr int RP_GROUPS = 32;
// Optional tail: the LAST CTA to finish (ticket) applies AdamW to the flat parameter buffer the gradient tensors are views
// of -- gm_adamw_step_ex's arithmetic, one launch less at the end of a single-GPU training step.  params == NULL: none.
struct AdamTail {
    float *params; const float *grads; float *m, *v;
    int n;
    float lr, b1, b2, eps, wd, bc1, bc2_sqrt, gscale;
    StepStatus status;
    int *ticket;
};
__global__ void __launch_bounds__(RP_GROUPS * 32) k_reduce_partials(Segs segs, AdamTail tail) {
    if (tail.params) pdl_wait(); else pdl_sync();        // with the tail: no early trigger (what follows may read weights up front)
    __shared__ float sm[RP_GROUPS][32];
    __shared__ int s_last;
    const Seg sg = segs.s[blockIdx.y];
    const int total = sg.rows * sg.cols;
    const int i = blockIdx.x * 32 + (threadIdx.x & 31);
    if (sg.dst && (int)blockIdx.x * 32 < total) {         // CTA-uniform
        const int g = threadIdx.x >> 5;
        float acc = 0.f;
        int r = 0, c = 0;
        if (i < total) {
            r = i / sg.cols; c = i - r * sg.cols;
            const float *p = sg.part + sg.off + r * sg.src_ld + c;
#pragma unroll 8
            for (int b = g; b < sg.nblk; b += RP_GROUPS) acc += p[(size_t)b * GP_STRIDE];
        }
        sm[g][threadIdx.x & 31] = acc;
        __syncthreads();
        if (g == 0 && i < total) {
            float t = 0.f;
#pragma unroll
            for (int k = 0; k < RP_GROUPS; ++k) t += sm[k][threadIdx.x];
            sg.dst[r * sg.dst_ld + c] = t;
        }
    }
    if (!tail.params) return;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const int n_ctas = (int)(gridDim.x * gridDim.y);
        const int last = atomicAdd(tail.ticket, 1) == n_ctas - 1;
        if (last) *tail.ticket = 0;
        s_last = last;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (threadIdx.x == 0) tail.status.publish();
    for (int k = threadIdx.x; k < tail.n; k += RP_GROUPS * 32) {
        const float gi = __ldcg(tail.grads + k) * tail.gscale;
        float pi = tail.params[k] * (1.f - tail.lr * tail.wd);
        const float mi = tail.b1 * tail.m[k] + (1.f - tail.b1) * gi;
        const float vi = tail.b2 * tail.v[k] + (1.f - tail.b2) * gi * gi;
        const float denom = sqrtf(vi) / tail.bc2_sqrt + tail.eps;
        pi -= (tail.lr / tail.bc1) * (mi / denom);
        tail.params[k] = pi; tail.m[k] = mi; tail.v[k] = vi;
    }
}

// Last CTA: fold the per-CTA (sum G, sum G*xhat) partials in a fixed order.  blockDim.x >= NODE_TILE: 16 lane
// groups of 8 lanes (two float4 = 8 of the 64 sums each) stream CTAs g, g+16, ... with independent loads,
// then 64 threads fold the 16 group results in group order.  scratch >= 1024 floats.
__device__ __forceinline__ void bn_sums_finalize(const BnSumOut &o, float *scratch) {
    const int t = threadIdx.x, sub = t & 7, grp = t >> 3;
    float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
#pragma unroll 4
    for (int b = grp; b < (int)gridDim.x && t < NODE_TILE; b += NODE_TILE / 8) {
        const float4 *p = reinterpret_cast<const float4 *>(o.part + (size_t)b * 64) + 2 * sub;
        const float4 u = __ldcg(p), v = __ldcg(p + 1);
        a0.x += u.x; a0.y += u.y; a0.z += u.z; a0.w += u.w;
        a1.x += v.x; a1.y += v.y; a1.z += v.z; a1.w += v.w;
    }
    if (t < NODE_TILE) {
        float *d = scratch + grp * 64 + 8 * sub;
        d[0] = a0.x; d[1] = a0.y; d[2] = a0.z; d[3] = a0.w; d[4] = a1.x; d[5] = a1.y; d[6] = a1.z; d[7] = a1.w;
    }
    __syncthreads();
    if (t < 64) {
        float tot = 0.f;
#pragma unroll
        for (int g = 0; g < NODE_TILE / 8; ++g) tot += scratch[g * 64 + t];
        o.sum[t] = tot;
        if (t < 32) { if (o.dbeta) o.dbeta[t] = tot; }
        else if (o.dgamma) o.dgamma[t - 32] = tot;
    }
    __syncthreads();
}

// ----------------------------------------------------------------------------------------- head backward
// partial layout: dW1[16][32] @0, db1[16] @512, dW2[16] @528, db2 @544
__global__ void __launch_bounds__(NODE_TILE, GM_BWD_OCC_HEAD)
k_head_bwd(const float *__restrict__ emb, const float *__restrict__ out, const float *__restrict__ grad_out,
           const int *__restrict__ perm, int N, const float *__restrict__ w1, const float *__restrict__ b1,
           const float *__restrict__ w2, int apply_sigmoid, float *__restrict__ G, float *__restrict__ gpart,
           BnSumOut bo) {
    __shared__ __align__(16) float s_w1[16 * H];
    __shared__ float s_b1[16];
    __shared__ float s_w2[16];
    __shared__ float s_mean[H], s_inv[H];
    __shared__ __align__(16) float tA[NODE_TILE * GLD];   // dh[0:16] | go*hact[16:32] | go at [32]   then  G row
    __shared__ __align__(16) float tB[NODE_TILE * GLD];   // emb row                                   then  G*xhat row
    __shared__ __align__(16) float scratch[1024];
    const int t = threadIdx.x;
    for (int i = t; i < 16 * H; i += NODE_TILE) s_w1[i] = w1[i];
    if (t < 16) { s_b1[t] = b1[t]; s_w2[t] = w2[t]; }
    if (t < H) { s_mean[t] = bo.stat[t]; s_inv[t] = bo.stat[2 * H + t]; }
    pdl_sync();                                  // above: parameters and forward statistics (many kernels back)
    __syncthreads();
    Outer<16, H> oW1; oW1.zero();
    ColSum<GLD> cs; cs.zero();   // columns 0..15 = db1, 16..31 = dW2, 32 = db2
    ColSum<H> sg, sgx; sg.zero(); sgx.zero();
    const int tiles = (N + NODE_TILE - 1) / NODE_TILE;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const int node0 = tile * NODE_TILE, rows = min(NODE_TILE, N - node0);
        float gx[H];
        if (t < rows) {
            const int n = node0 + t;
            const int src = perm ? perm[n] : n;
            float xr[H], hdn[16], dh[16];
            load_row32(emb + (size_t)n * H, xr);
            dense_smem<H, 16>(s_w1, s_b1, xr, hdn);
            const float w = out[src];
            const float go = grad_out[src] * (apply_sigmoid ? w * (1.f - w) : 1.f);
            float *ra = tA + t * GLD;
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                const float ha = elu(hdn[k]);
                dh[k] = go * s_w2[k] * elu_grad_from_out(ha);
                ra[k] = dh[k];
                ra[16 + k] = go * ha;
            }
            ra[32] = go; ra[33] = 0.f; ra[34] = 0.f; ra[35] = 0.f;
            dense_t_smem<H, 16>(s_w1, dh, gx);
            store_row<H>(G + (size_t)n * H, gx);
            store_row<H>(tB + t * GLD, xr);
        }
        __syncthreads();
        oW1.accumulate(tA, GLD, tB, GLD, rows);
        cs.accumulate(tA, GLD, rows);
        __syncthreads();
        if (t < rows) {   // BatchNorm sums of the layer below: G and G * xhat
            float raw[H];
            load_row32(bo.raw + (size_t)(node0 + t) * H, raw);
            store_row<H>(tA + t * GLD, gx);
#pragma unroll
            for (int k = 0; k < H; ++k) raw[k] = gx[k] * ((raw[k] - s_mean[k]) * s_inv[k]);
            store_row<H>(tB + t * GLD, raw);
        }
        __syncthreads();
        sg.accumulate(tA, GLD, rows);
        sgx.accumulate(tB, GLD, rows);
        __syncthreads();
    }
    // accumulators parked in the dead tiles, one barrier, fixed-order folds
    oW1.stage(tA); cs.stage(tB); sg.stage(tB + 128); sgx.stage(tB + 256);
    __syncthreads();
    float *dst = gpart + (size_t)blockIdx.x * GP_STRIDE;
    oW1.fold_staged(tA, dst);
    cs.fold_staged(tB, dst + 512, 33);
    float *bp = bo.part + (size_t)blockIdx.x * 64;
    sg.fold_staged(tB + 128, bp, H);
    sgx.fold_staged(tB + 256, bp + H, H);
    if (last_cta_done(bo.counter)) bn_sums_finalize(bo, scratch);
}

// ----------------------------------------------------------------------------------------- max-aggregation routing
// S[j][c] = sum over targets i that list j as a source of [arg(i,c) == j] * d(conv)[i][c].
// Warp per source j, half-warp per arg row (int2 per lane), up to 16 loads (32 targets) in flight per lane,
// next row prefetched; G and
// raw are only touched where a channel matches (N*32 matches in total).  Targets are visited in CSR order
// and the two halves are folded as (half 0) + (half 1) => fixed summation order.
// Matching (target, channel) entries are sparse (32 per source on average): the G / raw loads are PREDICATED,
// not branched, so all of a batch's loads are issued back to back (one more round trip, not one per match).
// A user edge_index may list an edge (j -> i) more than once (e.g. the dz variant: torch.cat of two radius lists,
// evaluate.py:91-96): the scatter-max of the forward routes a (target, channel) gradient to ONE edge, so repeated
// entries of a (sorted) transposed row are skipped -- a duplicate is marked -1 when the row is loaded.
__device__ __forceinline__ void scatter_consume(float2 &s, const int2 (&a)[8], const int (&i)[8], int j, int sub,
                                                const float *s_coef, const float2 *__restrict__ G2,
                                                const float2 *__restrict__ R2) {
    float2 g[8], r[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
        g[u] = make_float2(0.f, 0.f); r[u] = g[u];
        if (a[u].x == j || a[u].y == j) {          // j >= 0 never equals the -1 filler
            g[u] = __ldg(G2 + (size_t)i[u] * (H / 2) + sub);
            r[u] = __ldg(R2 + (size_t)i[u] * (H / 2) + sub);
        }
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
        const float dx = bn_bwd_one(s_coef, 2 * sub + 0, g[u].x, r[u].x);
        const float dy = bn_bwd_one(s_coef, 2 * sub + 1, g[u].y, r[u].y);
        s.x += a[u].x == j ? dx : 0.f;
        s.y += a[u].y == j ? dy : 0.f;
    }
}

__global__ void __launch_bounds__(256) k_scatter_s(const float *__restrict__ G, const float *__restrict__ raw, BnBwd bn,
                                                   const int *__restrict__ arg, const int *__restrict__ rowptr_s,
                                                   const int *__restrict__ col_s, const int *__restrict__ rowptr_tt,
                                                   const int *__restrict__ col_tt, const int *__restrict__ use_t, int N,
                                                   float *__restrict__ S) {
    __shared__ float s_coef[5 * H];
    bn_bwd_coeffs(bn, s_coef);
    __syncthreads();
    const bool tr = use_t ? (*use_t != 0) : true;
    const int *__restrict__ rowptr_t = tr ? rowptr_tt : rowptr_s;
    const int *__restrict__ col_t = tr ? col_tt : col_s;
    const int lane = threadIdx.x & 31, sub = lane & 15, half = lane >> 4;
    const int stride = (gridDim.x * blockDim.x) >> 5;
    const int2 *__restrict__ arg2 = reinterpret_cast<const int2 *>(arg);
    const float2 *__restrict__ G2 = reinterpret_cast<const float2 *>(G);
    const float2 *__restrict__ R2 = reinterpret_cast<const float2 *>(raw);
    const int2 NONE = make_int2(-1, -1);
    int j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int beg = 0, end = 0, il = -1;
    if (j < N) {
        beg = __ldg(rowptr_t + j); end = __ldg(rowptr_t + j + 1);
        il = lane < end - beg ? __ldg(col_t + beg + lane) : -1;
        if (lane > 0 && lane < end - beg && il == __ldg(col_t + beg + lane - 1)) il = -1;
    }
    while (j < N) {
        const int jn = j + stride;
        int nbeg = 0, nend = 0, inext = -1;
        if (jn < N) { nbeg = __ldg(rowptr_t + jn); nend = __ldg(rowptr_t + jn + 1); }
        float2 s = make_float2(0.f, 0.f);
        for (int e0 = beg; e0 < end; e0 += 32) {
            const int nb = min(32, end - e0);
            if (e0 > beg) {
                il = lane < nb ? __ldg(col_t + e0 + lane) : -1;
                if (lane < nb && il == __ldg(col_t + e0 + lane - 1)) il = -1;
            }
            int ia[8], ib[8];
            int2 aa[8], ab[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) ia[u] = __shfl_sync(0xffffffffu, il, 2 * u + half);
#pragma unroll
            for (int u = 0; u < 8; ++u) aa[u] = ia[u] >= 0 ? __ldg(arg2 + (size_t)ia[u] * (H / 2) + sub) : NONE;
            if (nb > 16) {
#pragma unroll
                for (int u = 0; u < 8; ++u) ib[u] = __shfl_sync(0xffffffffu, il, 16 + 2 * u + half);
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    ab[u] = ib[u] >= 0 ? __ldg(arg2 + (size_t)ib[u] * (H / 2) + sub) : NONE;
            }
            if (e0 == beg && jn < N) {
                inext = lane < nend - nbeg ? __ldg(col_t + nbeg + lane) : -1;
                if (lane > 0 && lane < nend - nbeg && inext == __ldg(col_t + nbeg + lane - 1)) inext = -1;
            }
            scatter_consume(s, aa, ia, j, sub, s_coef, G2, R2);
            if (nb > 16) scatter_consume(s, ab, ib, j, sub, s_coef, G2, R2);
        }
        if (end == beg && jn < N) {
            inext = lane < nend - nbeg ? __ldg(col_t + nbeg + lane) : -1;
            if (lane > 0 && lane < nend - nbeg && inext == __ldg(col_t + nbeg + lane - 1)) inext = -1;
        }
        const float ox = __shfl_xor_sync(0xffffffffu, s.x, 16), oy = __shfl_xor_sync(0xffffffffu, s.y, 16);
        if (half == 0) reinterpret_cast<float2 *>(S)[(size_t)j * (H / 2) + sub] = make_float2(s.x + ox, s.y + oy);
        j = jn; beg = nbeg; end = nend; il = inext;
    }
}

// ----------------------------------------------------------------------------------------- EdgeConv node-level backward
// partial layout: dW1[32][32] @0, dW2[32][32] @1024, db[32] @2048
// Same stage on the register-blocked tile product (common.cuh): 256 threads per 128-node tile.  Cooperative,
// coalesced float4 loads build the tiles d(conv) = BatchNorm backward of G, S and x in shared memory; the two
// 32x32 transposed products run as tile_gemm with the accumulators initialised from G (4 nodes x 4 channels per
// thread) and the new G leaves from registers as whole 128-byte rows; the parameter-gradient outer products and
// the BatchNorm sums of the layer below read the same tiles.  [c][k] weights are K-major for the transposed product.
constexpr int CB_THREADS = 256;
__global__ void __launch_bounds__(CB_THREADS, GM_BWD_OCC_CONV)
k_conv_bwd_node2(const float *__restrict__ x, const float *__restrict__ raw, BnBwd bn, const float *__restrict__ S, int N,
                 const float *__restrict__ W, float *__restrict__ G, float *__restrict__ gpart, BnSumOut bo) {
    extern __shared__ __align__(16) float smem[];
    float *s_wd = smem;                  // W1 - W2, [c][k]
    float *s_w2 = s_wd + H * H;
    float *tA = s_w2 + H * H;            // d(conv)            then  G_new
    float *tB = tA + NODE_TILE * GLD;    // x                  then  G_new * xhat(below)
    float *tC = tB + NODE_TILE * GLD;    // S                  then  S - d(conv)
    float *s_coef = tC + NODE_TILE * GLD;   // [5][32]
    float *s_mean = s_coef + 5 * H;      // BatchNorm below: mean, invstd
    float *s_inv = s_mean + H;
    const int t = threadIdx.x, ng = t >> 3, og = t & 7;
    for (int i = t; i < H * H; i += CB_THREADS) {
        const int c = i >> 5, k = i & 31;
        const float a = W[c * CONV_IN + k], b = W[c * CONV_IN + H + k];
        s_wd[i] = a - b; s_w2[i] = b;
    }
    pdl_sync();                                  // weights above: parameters; the BatchNorm sums are the previous kernel's
    bn_bwd_coeffs(bn, s_coef);
    if (t < H) { s_mean[t] = bo.stat[t]; s_inv[t] = bo.stat[2 * H + t]; }
    __syncthreads();
    Outer<H, H, CB_THREADS> o1, o2; o1.zero(); o2.zero();
    ColSum<H, CB_THREADS> cb, sg, sgx; cb.zero(); sg.zero(); sgx.zero();
    const int tiles = (N + NODE_TILE - 1) / NODE_TILE;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const int node0 = tile * NODE_TILE, rows = min(NODE_TILE, N - node0);
        const float4 *G4 = reinterpret_cast<const float4 *>(G) + (size_t)node0 * (H / 4);
        const float4 *R4 = reinterpret_cast<const float4 *>(raw) + (size_t)node0 * (H / 4);
        const float4 *S4 = reinterpret_cast<const float4 *>(S) + (size_t)node0 * (H / 4);
        const float4 *X4 = reinterpret_cast<const float4 *>(x) + (size_t)node0 * (H / 4);
#pragma unroll
        for (int it = 0; it < NODE_TILE * (H / 4) / CB_THREADS; ++it) {
            const int idx = it * CB_THREADS + t, r = idx >> 3, c4 = (idx & 7) * 4;
            float4 dc = make_float4(0.f, 0.f, 0.f, 0.f), sv = dc, xv = dc;
            if (r < rows) {
                const float4 g = G4[idx], rw = R4[idx];
                sv = S4[idx]; xv = X4[idx];
                dc.x = bn_bwd_one(s_coef, c4 + 0, g.x, rw.x); dc.y = bn_bwd_one(s_coef, c4 + 1, g.y, rw.y);
                dc.z = bn_bwd_one(s_coef, c4 + 2, g.z, rw.z); dc.w = bn_bwd_one(s_coef, c4 + 3, g.w, rw.w);
            }
            *reinterpret_cast<float4 *>(tA + r * GLD + c4) = dc;
            *reinterpret_cast<float4 *>(tC + r * GLD + c4) = sv;
            *reinterpret_cast<float4 *>(tB + r * GLD + c4) = xv;
        }
        __syncthreads();
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float4 g = ng + 32 * i < rows ? G4[(ng + 32 * i) * (H / 4) + og] : make_float4(0.f, 0.f, 0.f, 0.f);
            acc[i][0] = g.x; acc[i][1] = g.y; acc[i][2] = g.z; acc[i][3] = g.w;
        }
        tile_gemm<H, H, GLD>(tA, s_wd, ng, og, acc);       // += (W1 - W2)^T d(conv)
        tile_gemm<H, H, GLD>(tC, s_w2, ng, og, acc);       // += W2^T S
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (ng + 32 * i < rows)
                reinterpret_cast<float4 *>(G)[(size_t)(node0 + ng + 32 * i) * (H / 4) + og] =
                    make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        __syncthreads();                                   // products done with tA / tC
#pragma unroll
        for (int it = 0; it < NODE_TILE * (H / 4) / CB_THREADS; ++it) {
            const int idx = it * CB_THREADS + t, r = idx >> 3, c4 = (idx & 7) * 4;
            const float4 dc = *reinterpret_cast<const float4 *>(tA + r * GLD + c4);
            float4 sv = *reinterpret_cast<const float4 *>(tC + r * GLD + c4);
            sv.x -= dc.x; sv.y -= dc.y; sv.z -= dc.z; sv.w -= dc.w;
            *reinterpret_cast<float4 *>(tC + r * GLD + c4) = sv;
        }
        __syncthreads();
        o1.accumulate(tA, GLD, tB, GLD, rows);
        o2.accumulate(tC, GLD, tB, GLD, rows);
        cb.accumulate(tA, GLD, rows);
        __syncthreads();
        // BatchNorm sums of the layer below (bn_all for layer 0): tiles of G_new and G_new * xhat
        const float4 *RB4 = reinterpret_cast<const float4 *>(bo.raw) + (size_t)node0 * (H / 4);
        const float4 m4 = reinterpret_cast<const float4 *>(s_mean)[og], i4 = reinterpret_cast<const float4 *>(s_inv)[og];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int r = ng + 32 * i;
            float4 gx = make_float4(0.f, 0.f, 0.f, 0.f);
            if (r < rows) {
                const float4 rb = RB4[r * (H / 4) + og];
                gx.x = acc[i][0] * ((rb.x - m4.x) * i4.x); gx.y = acc[i][1] * ((rb.y - m4.y) * i4.y);
                gx.z = acc[i][2] * ((rb.z - m4.z) * i4.z); gx.w = acc[i][3] * ((rb.w - m4.w) * i4.w);
            }
            *reinterpret_cast<float4 *>(tA + r * GLD + 4 * og) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
            *reinterpret_cast<float4 *>(tB + r * GLD + 4 * og) = gx;
        }
        __syncthreads();
        sg.accumulate(tA, GLD, rows);
        sgx.accumulate(tB, GLD, rows);
        __syncthreads();
    }
    // all accumulators parked in the dead tiles (tA | tB | tC contiguous), one barrier, fixed-order folds
    float *st1 = tA, *st2 = st1 + 4 * 1024, *stc = st2 + 4 * 1024;
    o1.stage(st1); o2.stage(st2); cb.stage(stc); sg.stage(stc + 256); sgx.stage(stc + 512);
    __syncthreads();
    float *dst = gpart + (size_t)blockIdx.x * GP_STRIDE;
    o1.fold_staged(st1, dst);
    o2.fold_staged(st2, dst + 1024);
    cb.fold_staged(stc, dst + 2048, H);
    float *bp = bo.part + (size_t)blockIdx.x * 64;
    sg.fold_staged(stc + 256, bp, H);
    sgx.fold_staged(stc + 512, bp + H, H);
    if (last_cta_done(bo.counter)) bn_sums_finalize(bo, tA);
}

// ----------------------------------------------------------------------------------------- encoder backward
// partial layout: dWa[32][32] @0, dba[32] @1024, dWc[16][8] @1056, dbc[16] @1184, dWk[16][24] @1200,
//                 dbk[16] @1584, d(tables)[18][8] @1600 (charge rows 0-2, pdgid 3-9, pv 10-17)
struct EncBwdArgs {
    const float *x_cont; int x_cont_stride;
    const long long *x_cat_i64;
    const float *x_cat_f32; int x_cat_stride;
    const int *perm;
    int N, CD;
    const float *cont_w, *cont_b, *cat_w, *enc_w, *lut, *e_chrg, *e_pdg, *e_pv;
    const float *enc;      // [N,32] forward ELU(encode_all) output (pre-BatchNorm)
    const float *G;        // [N,32] gradient w.r.t. emb_0 = bn_all(enc)
    BnBwd bn;
    float *gpart;
};

// 256 threads per 128-node tile.  Per tile, four barrier intervals:
//   A  every thread: coalesced float4 loads of enc / G -> dz = BN'(G)·ELU'(enc) into tA; threads < 128: one node each,
//      rebuild u = [LUT row | ELU(W_c x + b)] into tB (x_c and the three table indices stay in registers)
//      [threads >= 112 meanwhile fold the PREVIOUS tile's d(table values) (tD) into the embedding-table accumulators]
//   B  du = dz·W_enc as a register-blocked tile product (4 nodes x 4 inputs per thread), times ELU'(u) read back from
//      tB; dW_enc += dzᵀu, db_enc += Σ dz
//   B' tA <- [d cat-pre | d cont-pre] from the product's registers, tB <- [x_c | the node's 24 table values]
//   C  dW_c, db_c, dW_k, db_k outer products / column sums; d(table values) = d cat-pre · W_k, (node, half) per thread -> tD
constexpr int EB_THREADS = 256;
constexpr int TD_LD = 28;                // d(table values) row: 24 floats + the node's packed table indices
__global__ void __launch_bounds__(EB_THREADS, GM_BWD_OCC_ENC) k_encode_bwd(EncBwdArgs a) {
    extern __shared__ __align__(16) float smem[];
    float *s_enc_w = smem;                       // 1024  [o][i]: K-major for du = dz·W
    float *s_cont_w = s_enc_w + H * H;           // 128   [16][8]
    float *s_cont_b = s_cont_w + 128;            // 16
    float *s_cat_w = s_cont_b + 16;              // 384   [16][24]: K-major for dv = dpk·W_k
    float *s_lut = s_cat_w + 384;                // 2688
    float *s_tab = s_lut + LUT_ROWS * 16;        // 144   [18][8]
    float *tA = s_tab + 144;                     // NODE_TILE*GLD
    float *tB = tA + NODE_TILE * GLD;            // NODE_TILE*GLD
    float *tD = tB + NODE_TILE * GLD;            // NODE_TILE*TD_LD
    float *scratch = tD + NODE_TILE * TD_LD;     // 1024
    float *s_coef = scratch + 1024;              // 160
    const int t = threadIdx.x, ng = t >> 3, og = t & 7;
    for (int i = t; i < H * H; i += EB_THREADS) s_enc_w[i] = a.enc_w[i];
    if (t < 128) { int o = t >> 3, k = t & 7; s_cont_w[t] = k < a.CD ? a.cont_w[o * a.CD + k] : 0.f; }
    if (t < 16) s_cont_b[t] = a.cont_b[t];
    for (int i = t; i < 384; i += EB_THREADS) s_cat_w[i] = a.cat_w[i];
    for (int i = t; i < LUT_ROWS * 16; i += EB_THREADS) s_lut[i] = a.lut[i];
    if (t < 144) s_tab[t] = t < 24 ? a.e_chrg[t] : (t < 80 ? a.e_pdg[t - 24] : a.e_pv[t - 80]);
    pdl_sync();                                  // above: parameters and the forward LUT; the BatchNorm sums are the previous kernel's
    bn_bwd_coeffs(a.bn, s_coef);
    __syncthreads();

    Outer<H, H, EB_THREADS> oA; oA.zero();
    Outer<16, 8, EB_THREADS> oC; oC.zero();
    Outer<16, 24, EB_THREADS> oK; oK.zero();
    ColSum<H, EB_THREADS> cA; cA.zero();
    ColSum<16, EB_THREADS> cC; cC.zero();
    ColSum<16, EB_THREADS> cK; cK.zero();
    float tab_acc = 0.f;
    // embedding-table accumulator of this thread: row r of the stacked 18-row table, column k
    const int tid = t - (EB_THREADS - 144);
    const int tr = tid >> 3, tk = tid & 7;
    const int tshift = tr < 3 ? 0 : (tr < 10 ? 8 : 16);
    const int twant = tr < 3 ? tr : (tr < 10 ? tr - 3 : tr - 10);
    const int tcol = tshift + tk;
    int prev_rows = 0;

    const int tiles = (a.N + NODE_TILE - 1) / NODE_TILE;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const int node0 = tile * NODE_TILE, rows = min(NODE_TILE, a.N - node0);
        // ---- A
        {
            const float4 *E4 = reinterpret_cast<const float4 *>(a.enc) + (size_t)node0 * (H / 4);
            const float4 *G4 = reinterpret_cast<const float4 *>(a.G) + (size_t)node0 * (H / 4);
#pragma unroll
            for (int it = 0; it < NODE_TILE * (H / 4) / EB_THREADS; ++it) {
                const int idx = it * EB_THREADS + t, r = idx >> 3, c4 = (idx & 7) * 4;
                float4 dz = make_float4(0.f, 0.f, 0.f, 0.f);
                if (r < rows) {
                    const float4 e = E4[idx], g = G4[idx];
                    dz.x = bn_bwd_one(s_coef, c4 + 0, g.x, e.x) * elu_grad_from_out(e.x);
                    dz.y = bn_bwd_one(s_coef, c4 + 1, g.y, e.y) * elu_grad_from_out(e.y);
                    dz.z = bn_bwd_one(s_coef, c4 + 2, g.z, e.z) * elu_grad_from_out(e.z);
                    dz.w = bn_bwd_one(s_coef, c4 + 3, g.w, e.w) * elu_grad_from_out(e.w);
                }
                *reinterpret_cast<float4 *>(tA + r * GLD + c4) = dz;
            }
        }
        float xc[8];
        int packed = 0;
        if (t < NODE_TILE) {
            float u[H];
#pragma unroll
            for (int k = 0; k < H; ++k) u[k] = 0.f;
#pragma unroll
            for (int k = 0; k < 8; ++k) xc[k] = 0.f;
            if (t < rows) {
                const size_t n = node0 + t;
                const size_t src = a.perm ? (size_t)a.perm[n] : n;
                const float *row = a.x_cont + src * a.x_cont_stride;
#pragma unroll
                for (int k = 0; k < 8; ++k) xc[k] = k < a.CD ? row[k] : 0.f;
                long long c0, c1, c2;
                if (a.x_cat_i64) { const long long *cr = a.x_cat_i64 + src * 3; c0 = cr[0]; c1 = cr[1]; c2 = cr[2]; }
                else { const float *cr = a.x_cat_f32 + src * a.x_cat_stride; c0 = (long long)cr[0]; c1 = (long long)cr[1]; c2 = (long long)cr[2]; }
                const long long lc = c1 + 1, lp = pdg_remap(c0), lv = c2;   // clamp exactly like the forward
                const int ic = (int)(lc < 0 ? 0 : (lc > 2 ? 2 : lc));
                const int ip = (int)(lp < 0 ? 0 : (lp > 6 ? 6 : lp));
                const int iv = (int)(lv < 0 ? 0 : (lv > 7 ? 7 : lv));
                packed = ic | (ip << 8) | (iv << 16);
                const int combo = (ic * 7 + ip) * 8 + iv;
                float h0[16];
#pragma unroll
                for (int k = 0; k < 16; ++k) u[k] = s_lut[combo * 16 + k];
                dense_smem<8, 16>(s_cont_w, s_cont_b, xc, h0);
#pragma unroll
                for (int k = 0; k < 16; ++k) u[16 + k] = elu(h0[k]);
            }
            store_row<H>(tB + t * GLD, u);
        }
        if (tid >= 0) {                          // the previous tile's table gradients (tD is not touched before interval C)
#pragma unroll 4
            for (int rr = 0; rr < prev_rows; ++rr)
                if (((__float_as_int(tD[rr * TD_LD + 24]) >> tshift) & 0xff) == twant) tab_acc += tD[rr * TD_LD + tcol];
        }
        __syncthreads();
        // ---- B
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
        tile_gemm<H, H, GLD>(tA, s_enc_w, ng, og, acc);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float4 uv = *reinterpret_cast<const float4 *>(tB + (ng + 32 * i) * GLD + 4 * og);
            acc[i][0] *= elu_grad_from_out(uv.x); acc[i][1] *= elu_grad_from_out(uv.y);
            acc[i][2] *= elu_grad_from_out(uv.z); acc[i][3] *= elu_grad_from_out(uv.w);
        }
        oA.accumulate(tA, GLD, tB, GLD, rows);
        cA.accumulate(tA, GLD, rows);
        __syncthreads();
        // ---- B'
#pragma unroll
        for (int i = 0; i < 4; ++i)
            *reinterpret_cast<float4 *>(tA + (ng + 32 * i) * GLD + 4 * og) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        if (t < NODE_TILE) {
            const int ic = packed & 0xff, ip = (packed >> 8) & 0xff, iv = (packed >> 16) & 0xff;
            float v[H];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                v[k] = xc[k];
                v[8 + k] = s_tab[ic * 8 + k]; v[16 + k] = s_tab[24 + ip * 8 + k]; v[24 + k] = s_tab[80 + iv * 8 + k];
            }
            store_row<H>(tB + t * GLD, v);
        }
        __syncthreads();
        // ---- C: tA = [d cat-pre (16) | d cont-pre (16)], tB = [x_c (8) | table values (24)]
        oC.accumulate(tA + 16, GLD, tB, GLD, rows);
        cC.accumulate(tA + 16, GLD, rows);
        oK.accumulate(tA, GLD, tB + 8, GLD, rows);
        cK.accumulate(tA, GLD, rows);
        {
            const int n = t >> 1, hf = t & 1;
            float dv[12];
#pragma unroll
            for (int j = 0; j < 12; ++j) dv[j] = 0.f;
            const float *dp = tA + n * GLD;
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                const float d = dp[k];
                const float4 *w4 = reinterpret_cast<const float4 *>(s_cat_w + k * 24 + 12 * hf);
#pragma unroll
                for (int q = 0; q < 3; ++q) {
                    const float4 w = w4[q];
                    dv[4 * q + 0] = fmaf(d, w.x, dv[4 * q + 0]); dv[4 * q + 1] = fmaf(d, w.y, dv[4 * q + 1]);
                    dv[4 * q + 2] = fmaf(d, w.z, dv[4 * q + 2]); dv[4 * q + 3] = fmaf(d, w.w, dv[4 * q + 3]);
                }
            }
            float4 *d4 = reinterpret_cast<float4 *>(tD + n * TD_LD + 12 * hf);
#pragma unroll
            for (int q = 0; q < 3; ++q) d4[q] = make_float4(dv[4 * q], dv[4 * q + 1], dv[4 * q + 2], dv[4 * q + 3]);
            if (t < NODE_TILE) tD[t * TD_LD + 24] = __int_as_float(packed);
        }
        prev_rows = rows;
        __syncthreads();
    }
    if (tid >= 0) {
#pragma unroll 4
        for (int rr = 0; rr < prev_rows; ++rr)
            if (((__float_as_int(tD[rr * TD_LD + 24]) >> tshift) & 0xff) == twant) tab_acc += tD[rr * TD_LD + tcol];
    }
    __syncthreads();
    // every accumulator parked in the (now dead) tiles, one barrier, one fixed-order fold: tA | tB | tD are contiguous
    static_assert(2 * NODE_TILE * GLD + NODE_TILE * TD_LD >= 4 * 1024 + 32 * 128 + 10 * 384 + 3 * 256, "staging pool");
    float *st_oA = tA, *st_oC = st_oA + 4 * 1024, *st_oK = st_oC + 32 * 128, *st_cA = st_oK + 10 * 384,
          *st_cC = st_cA + 256, *st_cK = st_cC + 256;
    oA.stage(st_oA); oC.stage(st_oC); oK.stage(st_oK); cA.stage(st_cA); cC.stage(st_cC); cK.stage(st_cK);
    __syncthreads();
    float *dst = a.gpart + (size_t)blockIdx.x * GP_STRIDE;
    oA.fold_staged(st_oA, dst);
    cA.fold_staged(st_cA, dst + 1024, H);
    oC.fold_staged(st_oC, dst + 1056);
    cC.fold_staged(st_cC, dst + 1184, 16);
    oK.fold_staged(st_oK, dst + 1200);
    cK.fold_staged(st_cK, dst + 1584, 16);
    if (tid >= 0) dst[1600 + tid] = tab_acc;
}

constexpr size_t CONV_BWD_SMEM = (size_t)(2 * H * H + 3 * NODE_TILE * GLD + 5 * H + 2 * H) * 4;
constexpr size_t ENC_BWD_SMEM =
    (size_t)(H * H + 128 + 16 + 384 + LUT_ROWS * 16 + 144 + 2 * NODE_TILE * GLD + NODE_TILE * TD_LD + 1024 + 5 * H) * 4;

}  // namespace

int check_params(const GmParams *p);

}  // namespace gm

using namespace gm;

extern "C" {

size_t gm_net_bwd_ws_bytes(int N, int conv_depth) {
    (void)conv_depth;
    return bwd_ws_layout(N < 0 ? 0 : N, nullptr, nullptr);
}

}  // extern "C"

// Shared body of gm_net_backward (explicit CSR) and gm_net_backward_geo (implicit radius graph: geo_ws != NULL).
static int net_backward_impl(const GmParams *p, const GmParams *grads, const float *x_cont, int x_cont_stride,
                             const int64_t *x_cat_i64, const float *x_cat_f32, int x_cat_stride, const int *rowptr,
                             const int *col, const int *rowptr_t, const int *col_t, const int *use_transposed_flag,
                             const void *geo_ws, int B, int N, const int *node_perm, int training, int apply_sigmoid,
                             int aggr, const void *fwd_ws, size_t fwd_ws_bytes, void *bwd_ws, size_t bwd_ws_bytes,
                             const float *out, const float *grad_out, void *stream, const KnnLayers *kl = nullptr,
                             const GmAdamW *opt = nullptr) {
    GM_TRY(check_params(p));
    if (!grads || N < 0 || !fwd_ws || !bwd_ws) return GM_ERR_INVALID_ARGUMENT;
    if (aggr != GM_AGGR_MAX) return GM_ERR_UNSUPPORTED;
    if (N > 0 && (!x_cont || (!x_cat_i64 && !x_cat_f32) || (!geo_ws && !kl && (!rowptr_t || !col_t)) || !out || !grad_out))
        return GM_ERR_INVALID_ARGUMENT;
    const int depth = p->conv_depth;
    if (fwd_ws_bytes < net_ws_layout(N, depth, 1, nullptr, nullptr)) return GM_ERR_CAPACITY;
    if (bwd_ws_bytes < bwd_ws_layout(N, nullptr, nullptr)) return GM_ERR_CAPACITY;
    if (N == 0) return GM_OK;  // caller zero-fills gradients of an empty batch
    cudaStream_t st = as_stream(stream);
    NetWs f;
    net_ws_layout(N, depth, 1, (char *)fwd_ws, &f);
    BwdWs w;
    bwd_ws_layout(N, (char *)bwd_ws, &w);
    const int tiles = (N + NODE_TILE - 1) / NODE_TILE;
    const int ngrid_head = persistent_grid(tiles, GM_BWD_OCC_HEAD);   // thread per node
    const int ngrid_enc = persistent_grid(tiles, GM_BWD_OCC_ENC);
    const int ngrid_conv = persistent_grid(tiles, GM_BWD_OCC_CONV);   // k_conv_bwd_node2 (tile products)
    const float inv_n = training ? 1.f / (float)N : 0.f;
    // the tickets reset themselves; the memset (a node that breaks the dependent-launch chain) only guards a workspace
    // used for the first time -- gm_net_backward_geo_adamw's contract makes the caller hand over zeroed tickets instead
    if (!opt) GM_CUDA_TRY(cudaMemsetAsync(w.counters, 0, NUM_COUNTERS * sizeof(int), st));

    auto gpart = [&](int stage) { return w.gpart + (size_t)stage * MAX_BWD_CTAS * GP_STRIDE; };
    // BatchNorm index b: 0 = bn_all (raw = enc), l+1 = conv_bn[l] (raw = conv[l])
    auto bn_raw = [&](int b) { return b == 0 ? (const float *)f.enc : (const float *)f.conv[b - 1]; };
    auto bn_stat = [&](int b) { return (const float *)(f.bnstat + (size_t)b * BNSTAT_STRIDE); };
    auto bn_sum = [&](int b) { return w.bnsum + (size_t)b * 64; };
    auto bn_grads = [&](int b) -> const GmBatchNorm & { return b == 0 ? grads->bn_all : grads->conv_bn[b - 1]; };
    auto sum_out = [&](int b) {
        return BnSumOut{bn_raw(b), bn_stat(b), w.bnpart, bn_sum(b), bn_grads(b).weight, bn_grads(b).bias,
                        w.counters + b};
    };
    auto bn_in = [&](int b) { return BnBwd{bn_stat(b), bn_sum(b), inv_n}; };

    Segs segs{};
    auto add_seg = [&](int stage, float *dst, int dst_ld, int off, int rows, int src_ld, int cols) {
        const int ngrid = stage == 0 ? ngrid_head : stage <= depth ? ngrid_conv : ngrid_enc;
        if (dst) segs.s[segs.n++] = Seg{gpart(stage), dst, dst_ld, off, rows, src_ld, cols, ngrid};
    };

    // ---- head (stage 0); also produces the BatchNorm sums of the topmost BatchNorm (index depth)
    {
        TimedRegion tr(GM_TIMER_HEAD_BWD, st);
        GM_LAUNCH_PDL(k_head_bwd, ngrid_head, NODE_TILE, 0, st, f.emb[depth], out, grad_out, node_perm, N, p->out1_w, p->out1_b,
                      p->out2_w, apply_sigmoid, w.G, gpart(0), sum_out(depth));
        GM_LAUNCH_CHECK();
    }
    add_seg(0, grads->out1_w, H, 0, 16, H, H);
    add_seg(0, grads->out1_b, 16, 512, 1, 16, 16);
    add_seg(0, grads->out2_w, 16, 528, 1, 16, 16);
    add_seg(0, grads->out2_b, 1, 544, 1, 1, 1);

    // ---- EdgeConv layers, last to first
    GM_CUDA_TRY(cudaFuncSetAttribute(k_conv_bwd_node2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CONV_BWD_SMEM));
    for (int l = depth - 1; l >= 0; --l) {
        const int stage = depth - l;
        {
            TimedRegion tr(GM_TIMER_SCATTER_S_BWD, st);
            if (geo_ws) {
                GM_TRY(geo_scatter_s(w.G, f.conv[l], bn_in(l + 1), f.arg[l], geo_ws, N, B, w.S, st));
            } else {
                static int occ = 0;
                if (!occ) occ = resident_ctas(k_scatter_s, 256, 0, 3);
                // per-layer k-NN graphs are asymmetric: always the layer's own transposed CSR
                k_scatter_s<<<persistent_grid(((long long)N * 32 + 255) / 256, occ), 256, 0, st>>>(
                    w.G, f.conv[l], bn_in(l + 1), f.arg[l], kl ? kl->rowptr[l] : rowptr, kl ? kl->col[l] : col,
                    kl ? kl->rowptr_t[l] : rowptr_t, kl ? kl->col_t[l] : col_t, kl ? nullptr : use_transposed_flag, N, w.S);
                GM_LAUNCH_CHECK();
            }
        }
        {
            TimedRegion tr(GM_TIMER_CONV_NODE_BWD, st);
            GM_LAUNCH_PDL(k_conv_bwd_node2, ngrid_conv, CB_THREADS, CONV_BWD_SMEM, st, f.emb[l], f.conv[l], bn_in(l + 1), w.S, N,
                          p->conv_w[l], w.G, gpart(stage), sum_out(l));
            GM_LAUNCH_CHECK();
        }
        add_seg(stage, grads->conv_w[l], CONV_IN, 0, H, H, H);
        add_seg(stage, grads->conv_w[l] ? grads->conv_w[l] + H : nullptr, CONV_IN, 1024, H, H, H);
        add_seg(stage, grads->conv_b[l], H, 2048, 1, H, H);
    }
    // ---- bn_all + encoder (stage depth+1)
    {
        const int stage = depth + 1;
        EncBwdArgs ea{};
        ea.x_cont = x_cont; ea.x_cont_stride = x_cont_stride; ea.x_cat_i64 = (const long long *)x_cat_i64;
        ea.x_cat_f32 = x_cat_f32; ea.x_cat_stride = x_cat_stride; ea.perm = node_perm; ea.N = N;
        ea.CD = p->continuous_dim;
        ea.cont_w = p->cont_w; ea.cont_b = p->cont_b; ea.cat_w = p->cat_w; ea.enc_w = p->enc_w; ea.lut = f.lut;
        ea.e_chrg = p->embed_charge; ea.e_pdg = p->embed_pdgid; ea.e_pv = p->embed_pv;
        ea.enc = f.enc; ea.G = w.G; ea.bn = bn_in(0); ea.gpart = gpart(stage);
        GM_CUDA_TRY(cudaFuncSetAttribute(k_encode_bwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ENC_BWD_SMEM));
        {
            TimedRegion tr(GM_TIMER_ENCODE_BWD, st);
            GM_LAUNCH_PDL(k_encode_bwd, ngrid_enc, EB_THREADS, ENC_BWD_SMEM, st, ea);
            GM_LAUNCH_CHECK();
        }
        const int CD = p->continuous_dim;
        add_seg(stage, grads->enc_w, H, 0, H, H, H);
        add_seg(stage, grads->enc_b, H, 1024, 1, H, H);
        add_seg(stage, grads->cont_w, CD, 1056, 16, 8, CD);
        add_seg(stage, grads->cont_b, 16, 1184, 1, 16, 16);
        add_seg(stage, grads->cat_w, 24, 1200, 16, 24, 24);
        add_seg(stage, grads->cat_b, 16, 1584, 1, 16, 16);
        add_seg(stage, grads->embed_charge, 8, 1600, 3, 8, 8);
        add_seg(stage, grads->embed_pdgid, 8, 1624, 7, 8, 8);
        add_seg(stage, grads->embed_pv, 8, 1680, 8, 8, 8);
    }
    {
        int max_total = 0;
        for (int k = 0; k < segs.n; ++k) max_total = max(max_total, segs.s[k].rows * segs.s[k].cols);
        if (segs.n > 0) {
            dim3 grid((max_total + 31) / 32, segs.n);
            AdamTail tail{};
            if (opt) {
                const double bc1 = 1.0 - pow((double)opt->beta1, (double)opt->step);
                const double bc2 = 1.0 - pow((double)opt->beta2, (double)opt->step);
                tail = AdamTail{opt->params, opt->grads, opt->exp_avg, opt->exp_avg_sq, opt->n, opt->lr, opt->beta1, opt->beta2,
                                opt->eps, opt->weight_decay, (float)bc1, (float)sqrt(bc2), opt->grad_scale,
                                StepStatus{opt->err_flags_dev, opt->graph_stats_dev, opt->status_host, opt->step},
                                w.counters + NUM_COUNTERS - 1};
            }
            GM_LAUNCH_PDL(k_reduce_partials, grid, RP_GROUPS * 32, 0, st, segs, tail);
            GM_LAUNCH_CHECK();
        }
    }
    return GM_OK;
}

extern "C" {

int gm_net_backward(const GmParams *p, const GmParams *grads, const float *x_cont, int x_cont_stride,
                    const int64_t *x_cat_i64, const float *x_cat_f32, int x_cat_stride, const int *rowptr,
                    const int *col, const int *rowptr_t, const int *col_t, const int *use_transposed_flag, int N,
                    const int *node_perm, int training, int apply_sigmoid, int aggr, float eps, const void *fwd_ws,
                    size_t fwd_ws_bytes, void *bwd_ws, size_t bwd_ws_bytes, const float *out, const float *grad_out,
                    void *stream) {
    (void)eps;
    return net_backward_impl(p, grads, x_cont, x_cont_stride, x_cat_i64, x_cat_f32, x_cat_stride, rowptr, col, rowptr_t,
                             col_t, use_transposed_flag, nullptr, 0, N, node_perm, training, apply_sigmoid, aggr, fwd_ws,
                             fwd_ws_bytes, bwd_ws, bwd_ws_bytes, out, grad_out, stream);
}

int gm_net_backward_knn(const GmParams *p, const GmParams *grads, const float *x_cont, int x_cont_stride,
                        const int64_t *x_cat_i64, const float *x_cat_f32, int x_cat_stride, int N, int k, int training,
                        int apply_sigmoid, int aggr, const void *fwd_ws, size_t fwd_ws_bytes, void *bwd_ws,
                        size_t bwd_ws_bytes, const void *knn_ws, size_t knn_ws_bytes, const float *out,
                        const float *grad_out, void *stream) {
    if (!knn_ws || !p) return GM_ERR_INVALID_ARGUMENT;
    if (knn_ws_bytes < knn_layers_layout(N, p->conv_depth, k, nullptr, nullptr)) return GM_ERR_CAPACITY;
    KnnLayers kl;
    knn_layers_layout(N, p->conv_depth, k, (char *)knn_ws, &kl);
    return net_backward_impl(p, grads, x_cont, x_cont_stride, x_cat_i64, x_cat_f32, x_cat_stride, nullptr, nullptr,
                             nullptr, nullptr, nullptr, nullptr, 0, N, nullptr, training, apply_sigmoid, aggr, fwd_ws,
                             fwd_ws_bytes, bwd_ws, bwd_ws_bytes, out, grad_out, stream, &kl);
}

int gm_net_backward_geo(const GmParams *p, const GmParams *grads, const float *x_cont, int x_cont_stride,
                        const int64_t *x_cat_i64, const float *x_cat_f32, int x_cat_stride, const void *geo_ws, int N,
                        int B, const int *node_perm, int training, int apply_sigmoid, int aggr, const void *fwd_ws,
                        size_t fwd_ws_bytes, void *bwd_ws, size_t bwd_ws_bytes, const float *out,
                        const float *grad_out, void *stream) {
    if (!geo_ws || B < 1 || !node_perm) return GM_ERR_INVALID_ARGUMENT;
    return net_backward_impl(p, grads, x_cont, x_cont_stride, x_cat_i64, x_cat_f32, x_cat_stride, nullptr, nullptr,
                             nullptr, nullptr, nullptr, geo_ws, B, N, node_perm, training, apply_sigmoid, aggr, fwd_ws,
                             fwd_ws_bytes, bwd_ws, bwd_ws_bytes, out, grad_out, stream);
}

int gm_net_backward_geo_adamw(const GmParams *p, const GmParams *grads, const float *x_cont, int x_cont_stride,
                              const int64_t *x_cat_i64, const float *x_cat_f32, int x_cat_stride, const void *geo_ws, int N,
                              int B, const int *node_perm, int training, int apply_sigmoid, int aggr, const void *fwd_ws,
                              size_t fwd_ws_bytes, void *bwd_ws, size_t bwd_ws_bytes, const float *out,
                              const float *grad_out, const GmAdamW *opt, void *stream) {
    if (!geo_ws || B < 1 || !node_perm || N < 1) return GM_ERR_INVALID_ARGUMENT;
    if (!opt || opt->n < 1 || opt->step < 1 || !opt->params || !opt->grads || !opt->exp_avg || !opt->exp_avg_sq)
        return GM_ERR_INVALID_ARGUMENT;
    return net_backward_impl(p, grads, x_cont, x_cont_stride, x_cat_i64, x_cat_f32, x_cat_stride, nullptr, nullptr,
                             nullptr, nullptr, nullptr, geo_ws, B, N, node_perm, training, apply_sigmoid, aggr, fwd_ws,
                             fwd_ws_bytes, bwd_ws, bwd_ws_bytes, out, grad_out, stream, nullptr, opt);
}

}  // extern "C"
