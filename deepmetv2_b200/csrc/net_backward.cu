// Backward of the GraphMETNetwork hot path (what loss.backward() runs through autograd in the
// reference, train.py:51): hand-written, atomics-free, fixed-order (bit-deterministic) kernels.
//
//   k_head_bwd        : d(out) -> d(emb_L), partial dW/db of output.{0,2}
//   k_bn_bwd_reduce   : per-channel sum(G), sum(G*xhat)                     (BatchNorm backward, pass 1)
//   k_bn_bwd_apply    : d(conv) = gamma*invstd*(G - mean(G) - xhat*mean(G*xhat)); writes d gamma, d beta
//   k_scatter_s       : S[j][c] = sum_{i: argmax_src(i,c) == j} d(conv)[i][c]  as a GATHER over the
//                       transposed CSR (max-aggregation routes each (target,channel) gradient to one edge)
//   k_conv_bwd_node   : G += (W1-W2)^T d(conv) + W2^T S ; partial dW1 = dconv (x) x, dW2 = (S-dconv) (x) x
//   k_encode_bwd      : BN_all'd gradient -> encode_all / embed_continuous / embed_categorical / 3 tables
//   k_reduce_partials : fixed-order sum of the per-CTA partials into the caller's gradient tensors
//
// Parameter-gradient GEMMs (dW = sum_n a_n (x) b_n) are split-K over node tiles: every CTA keeps 4x4
// register blocks of the outer product for its tiles, the per-CTA results are summed in CTA order.
#include "net_layout.cuh"

namespace gm {
namespace {

constexpr int GLD = 36;                 // float4-aligned padded tile row (conflict-free LDS.128 row-per-thread)
constexpr int GP_STRIDE = 2112;         // floats per CTA in the parameter-gradient partial buffer
constexpr int BNB_THREADS = 256;
constexpr int MAX_BWD_CTAS = 148 * 2;

struct BwdWs {
    float *G, *DC, *S;                  // [N,32] each
    float *bnpart;                      // [MAX_BWD_CTAS][64]
    float *gpart;                       // [MAX_BWD_CTAS][GP_STRIDE]
};

size_t bwd_ws_layout(int N, char *base, BwdWs *w) {
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return base ? base + o : nullptr; };
    const size_t nodes = (size_t)(N > 0 ? N : 1);
    BwdWs t;
    t.G = (float *)take(nodes * H * sizeof(float));
    t.DC = (float *)take(nodes * H * sizeof(float));
    t.S = (float *)take(nodes * H * sizeof(float));
    t.bnpart = (float *)take((size_t)MAX_BWD_CTAS * 64 * sizeof(float));
    t.gpart = (float *)take((size_t)MAX_BWD_CTAS * GP_STRIDE * sizeof(float));
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
template <int O, int I>
struct Outer {
    static constexpr int BO = O / 4, BI = I / 4, BLK = BO * BI;
    static constexpr int GROUPS = NODE_TILE / BLK;
    static constexpr int ROWS_PER = (NODE_TILE + GROUPS - 1) / GROUPS;
    static_assert(O % 4 == 0 && I % 4 == 0 && BLK <= NODE_TILE, "bad outer-product shape");
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
    // fold the row groups in order, write the [O][I] CTA partial; scratch >= O*I floats of smem
    __device__ __forceinline__ void flush(float *scratch, float *__restrict__ dst) {
        const int t = threadIdx.x;
        for (int i = t; i < O * I; i += NODE_TILE) scratch[i] = 0.f;
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
        for (int i = t; i < O * I; i += NODE_TILE) dst[i] = scratch[i];
        __syncthreads();
    }
};

// column sums of a tile: s[c] += sum_rows T[r][c], c < NC
template <int NC>
struct ColSum {
    static constexpr int GROUPS = NODE_TILE / NC;
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

struct Seg { float *dst; int dst_ld, off, rows, src_ld, cols; };
struct Segs { int n; Seg s[10]; };

// dst[r*dst_ld + c] = sum_b part[b][off + r*src_ld + c].  grid = (chunks of 32 outputs, segments);
// thread (c = t%32, g = t/32): lanes read 32 consecutive outputs (coalesced), the 8 groups split the
// CTA-partial axis (b = g, g+8, ...) and are folded in a fixed order => deterministic.
__global__ void __launch_bounds__(256) k_reduce_partials(const float *__restrict__ part, int nblk, int stride, Segs segs) {
    __shared__ float sm[8][32];
    const Seg sg = segs.s[blockIdx.y];
    const int total = sg.rows * sg.cols;
    const int i = blockIdx.x * 32 + (threadIdx.x & 31);
    if (!sg.dst || blockIdx.x * 32 >= total) return;
    const int g = threadIdx.x >> 5;
    float acc = 0.f;
    int r = 0, c = 0;
    if (i < total) {
        r = i / sg.cols; c = i - r * sg.cols;
        const float *p = part + sg.off + r * sg.src_ld + c;
#pragma unroll 4
        for (int b = g; b < nblk; b += 8) acc += p[(size_t)b * stride];
    }
    sm[g][threadIdx.x & 31] = acc;
    __syncthreads();
    if (g == 0 && i < total) {
        float t = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) t += sm[k][threadIdx.x];
        sg.dst[r * sg.dst_ld + c] = t;
    }
}

int launch_reduce(const float *part, int nblk, const Segs &segs, cudaStream_t st) {
    int max_total = 0;
    for (int k = 0; k < segs.n; ++k) max_total = max(max_total, segs.s[k].rows * segs.s[k].cols);
    dim3 grid((max_total + 31) / 32, segs.n);
    k_reduce_partials<<<grid, 256, 0, st>>>(part, nblk, GP_STRIDE, segs);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

// ----------------------------------------------------------------------------------------- head backward
// partial layout: dW1[16][32] @0, db1[16] @512, dW2[16] @528, db2 @544
__global__ void __launch_bounds__(NODE_TILE, 2)
k_head_bwd(const float *__restrict__ emb, const float *__restrict__ out, const float *__restrict__ grad_out, int N,
           const float *__restrict__ w1, const float *__restrict__ b1, const float *__restrict__ w2,
           int apply_sigmoid, float *__restrict__ G, float *__restrict__ gpart) {
    __shared__ __align__(16) float s_w1[16 * H];
    __shared__ float s_b1[16];
    __shared__ float s_w2[16];
    __shared__ __align__(16) float tA[NODE_TILE * GLD];   // dh[0:16] | go*hact[16:32] | go at [32]
    __shared__ __align__(16) float tB[NODE_TILE * GLD];   // emb row
    __shared__ float scratch[16 * H];
    const int t = threadIdx.x;
    for (int i = t; i < 16 * H; i += NODE_TILE) s_w1[i] = w1[i];
    if (t < 16) { s_b1[t] = b1[t]; s_w2[t] = w2[t]; }
    __syncthreads();
    Outer<16, H> oW1; oW1.zero();
    ColSum<GLD> cs; cs.zero();   // columns 0..15 = db1, 16..31 = dW2, 32 = db2
    const int tiles = (N + NODE_TILE - 1) / NODE_TILE;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const int node0 = tile * NODE_TILE, rows = min(NODE_TILE, N - node0);
        if (t < rows) {
            const int n = node0 + t;
            float xr[H], hdn[16], dh[16], gx[H];
            load_row32(emb + (size_t)n * H, xr);
            dense_smem<H, 16>(s_w1, s_b1, xr, hdn);
            const float w = out[n];
            const float go = grad_out[n] * (apply_sigmoid ? w * (1.f - w) : 1.f);
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
    }
    float *dst = gpart + (size_t)blockIdx.x * GP_STRIDE;
    oW1.flush(scratch, dst);
    cs.flush(scratch, dst + 512, 33);
}

// ----------------------------------------------------------------------------------------- BatchNorm backward
// stat: mean[32], scale[32] (= gamma*invstd), invstd[32], beta[32]
__global__ void __launch_bounds__(BNB_THREADS) k_bn_bwd_reduce(const float4 *__restrict__ G, const float4 *__restrict__ raw,
                                                               const float *__restrict__ stat, size_t n4,
                                                               float *__restrict__ part) {
    __shared__ float sm[BNB_THREADS][9];
    const int t = threadIdx.x;
    const int c = (t & 7) * 4;   // gridDim.x*blockDim.x is a multiple of 8 => a thread keeps its channels
    float mean[4], inv[4], sg[4] = {0.f, 0.f, 0.f, 0.f}, sgx[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int k = 0; k < 4; ++k) { mean[k] = stat[c + k]; inv[k] = stat[2 * H + c + k]; }
    for (size_t i = (size_t)blockIdx.x * BNB_THREADS + t; i < n4; i += (size_t)gridDim.x * BNB_THREADS) {
        const float4 g = G[i], r = raw[i];
        sg[0] += g.x; sgx[0] = fmaf(g.x, (r.x - mean[0]) * inv[0], sgx[0]);
        sg[1] += g.y; sgx[1] = fmaf(g.y, (r.y - mean[1]) * inv[1], sgx[1]);
        sg[2] += g.z; sgx[2] = fmaf(g.z, (r.z - mean[2]) * inv[2], sgx[2]);
        sg[3] += g.w; sgx[3] = fmaf(g.w, (r.w - mean[3]) * inv[3], sgx[3]);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) { sm[t][k] = sg[k]; sm[t][4 + k] = sgx[k]; }
    __syncthreads();
    if (t < 64) {  // t<32: sum G of channel t ; t>=32: sum G*xhat of channel t-32
        const int ch = t & 31, which = t >> 5;
        const int lane8 = ch >> 2, sub = ch & 3;
        float acc = 0.f;
        for (int k = lane8; k < BNB_THREADS; k += 8) acc += sm[k][which * 4 + sub];
        part[(size_t)blockIdx.x * 64 + t] = acc;
    }
}

__global__ void __launch_bounds__(BNB_THREADS) k_bn_bwd_apply(const float4 *__restrict__ G, const float4 *__restrict__ raw,
                                                              const float *__restrict__ stat, const float *__restrict__ part,
                                                              int nblk, int N, int training, float *__restrict__ dgamma,
                                                              float *__restrict__ dbeta, float4 *__restrict__ DC, size_t n4) {
    __shared__ float s_sum[64];
    __shared__ float s_stat[BNSTAT_STRIDE];
    const int t = threadIdx.x;
    if (t < 64) {
        float acc = 0.f;
        for (int b = 0; b < nblk; ++b) acc += part[(size_t)b * 64 + t];
        s_sum[t] = acc;
        if (blockIdx.x == 0) {
            if (t < 32) { if (dbeta) dbeta[t] = acc; }
            else if (dgamma) dgamma[t - 32] = acc;
        }
    }
    if (t < BNSTAT_STRIDE) s_stat[t] = stat[t];
    __syncthreads();
    const float inv_n = training ? 1.f / (float)N : 0.f;
    const int c = (t & 7) * 4;
    float mean[4], inv[4], scale[4], c1[4], c2[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        mean[k] = s_stat[c + k]; scale[k] = s_stat[H + c + k]; inv[k] = s_stat[2 * H + c + k];
        c1[k] = s_sum[c + k] * inv_n; c2[k] = s_sum[32 + c + k] * inv_n;
    }
    for (size_t i = (size_t)blockIdx.x * BNB_THREADS + t; i < n4; i += (size_t)gridDim.x * BNB_THREADS) {
        const float4 g = G[i], r = raw[i];
        float4 o;
        o.x = scale[0] * (g.x - c1[0] - (r.x - mean[0]) * inv[0] * c2[0]);
        o.y = scale[1] * (g.y - c1[1] - (r.y - mean[1]) * inv[1] * c2[1]);
        o.z = scale[2] * (g.z - c1[2] - (r.z - mean[2]) * inv[2] * c2[2]);
        o.w = scale[3] * (g.w - c1[3] - (r.w - mean[3]) * inv[3] * c2[3]);
        DC[i] = o;
    }
}

// ----------------------------------------------------------------------------------------- max-aggregation routing
// S[j][c] = sum over targets i that list j as a source of [arg(i,c) == j] * DC[i][c].
// Warp per source j, lane = channel; the DC load is predicated on the match so only live sectors move.
__global__ void __launch_bounds__(256) k_scatter_s(const float *__restrict__ DC, const int *__restrict__ arg,
                                                   const int *__restrict__ rowptr_s, const int *__restrict__ col_s,
                                                   const int *__restrict__ rowptr_tt, const int *__restrict__ col_tt,
                                                   const int *__restrict__ use_t, int N, float *__restrict__ S) {
    const bool tr = use_t ? (*use_t != 0) : true;
    const int *__restrict__ rowptr_t = tr ? rowptr_tt : rowptr_s;
    const int *__restrict__ col_t = tr ? col_tt : col_s;
    const int lane = threadIdx.x & 31;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    for (int j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; j < N; j += warps) {
        const int beg = rowptr_t[j], end = rowptr_t[j + 1];
        float s = 0.f;
        for (int e0 = beg; e0 < end; e0 += 32) {
            const int il = (e0 + lane < end) ? col_t[e0 + lane] : -1;
            const int nb = min(32, end - e0);
            for (int e = 0; e < nb; e += 4) {
                int i[4], a[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    i[u] = __shfl_sync(0xffffffffu, il, (e + u) & 31);
                    a[u] = (e + u < nb) ? arg[(size_t)i[u] * H + lane] : -1;
                }
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (a[u] == j) s += DC[(size_t)i[u] * H + lane];
            }
        }
        S[(size_t)j * H + lane] = s;
    }
}

// ----------------------------------------------------------------------------------------- EdgeConv node-level backward
// partial layout: dW1[32][32] @0, dW2[32][32] @1024, db[32] @2048
__global__ void __launch_bounds__(NODE_TILE, 2)
k_conv_bwd_node(const float *__restrict__ x, const float *__restrict__ DC, const float *__restrict__ S, int N,
                const float *__restrict__ W, float *__restrict__ G, float *__restrict__ gpart) {
    extern __shared__ __align__(16) float smem[];
    float *s_wd = smem;                  // W1 - W2, [c][k]
    float *s_w2 = s_wd + H * H;
    float *tA = s_w2 + H * H;
    float *tB = tA + NODE_TILE * GLD;
    float *tC = tB + NODE_TILE * GLD;
    const int t = threadIdx.x;
    for (int i = t; i < H * H; i += NODE_TILE) {
        const int c = i >> 5, k = i & 31;
        const float a = W[c * CONV_IN + k], b = W[c * CONV_IN + H + k];
        s_wd[i] = a - b; s_w2[i] = b;
    }
    __syncthreads();
    Outer<H, H> o1, o2; o1.zero(); o2.zero();
    ColSum<H> cb; cb.zero();
    const int tiles = (N + NODE_TILE - 1) / NODE_TILE;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const int node0 = tile * NODE_TILE, rows = min(NODE_TILE, N - node0);
        if (t < rows) {
            const size_t n = (size_t)(node0 + t) * H;
            float dc[H], s[H], tmp[H], g[H];
            load_row32(DC + n, dc);
            load_row32(S + n, s);
            load_row32(G + n, g);
            dense_t_smem<H, H>(s_wd, dc, tmp);
#pragma unroll
            for (int k = 0; k < H; ++k) g[k] += tmp[k];
            dense_t_smem<H, H>(s_w2, s, tmp);
#pragma unroll
            for (int k = 0; k < H; ++k) g[k] += tmp[k];
            store_row<H>(G + n, g);
            store_row<H>(tA + t * GLD, dc);
#pragma unroll
            for (int k = 0; k < H; ++k) s[k] -= dc[k];
            store_row<H>(tC + t * GLD, s);
            load_row32(x + n, tmp);
            store_row<H>(tB + t * GLD, tmp);
        }
        __syncthreads();
        o1.accumulate(tA, GLD, tB, GLD, rows);
        o2.accumulate(tC, GLD, tB, GLD, rows);
        cb.accumulate(tA, GLD, rows);
        __syncthreads();
    }
    float *dst = gpart + (size_t)blockIdx.x * GP_STRIDE;
    o1.flush(tA, dst);
    o2.flush(tA, dst + 1024);
    cb.flush(tA, dst + 2048, H);
}

// ----------------------------------------------------------------------------------------- encoder backward
// partial layout: dWa[32][32] @0, dba[32] @1024, dWc[16][8] @1056, dbc[16] @1184, dWk[16][24] @1200,
//                 dbk[16] @1584, d(tables)[18][8] @1600 (charge rows 0-2, pdgid 3-9, pv 10-17)
struct EncBwdArgs {
    const float *x_cont; int x_cont_stride;
    const long long *x_cat_i64;
    const float *x_cat_f32; int x_cat_stride;
    int N, CD;
    const float *cont_w, *cont_b, *cat_w, *enc_w, *lut, *e_chrg, *e_pdg, *e_pv;
    const float *enc;      // [N,32] forward ELU(encode_all) output
    const float *DA;       // [N,32] gradient w.r.t. enc (after BN_all backward)
    float *gpart;
};

__global__ void __launch_bounds__(NODE_TILE, 2) k_encode_bwd(EncBwdArgs a) {
    extern __shared__ __align__(16) float smem[];
    float *s_enc_w = smem;                       // 1024
    float *s_cont_w = s_enc_w + H * H;           // 128  [16][8]
    float *s_cont_b = s_cont_w + 128;            // 16
    float *s_cat_w = s_cont_b + 16;              // 384  [16][24]
    float *s_lut = s_cat_w + 384;                // 2688
    float *s_tab = s_lut + LUT_ROWS * 16;        // 144  [18][8]
    float *tA = s_tab + 144;                     // NODE_TILE*GLD
    float *tB = tA + NODE_TILE * GLD;            // NODE_TILE*GLD
    float *scratch = tB + NODE_TILE * GLD;       // 1024
    int *s_idx = reinterpret_cast<int *>(scratch + 1024);  // NODE_TILE
    const int t = threadIdx.x;
    for (int i = t; i < H * H; i += NODE_TILE) s_enc_w[i] = a.enc_w[i];
    for (int i = t; i < 128; i += NODE_TILE) { int o = i >> 3, k = i & 7; s_cont_w[i] = k < a.CD ? a.cont_w[o * a.CD + k] : 0.f; }
    if (t < 16) s_cont_b[t] = a.cont_b[t];
    for (int i = t; i < 384; i += NODE_TILE) s_cat_w[i] = a.cat_w[i];
    for (int i = t; i < LUT_ROWS * 16; i += NODE_TILE) s_lut[i] = a.lut[i];
    for (int i = t; i < 144; i += NODE_TILE)
        s_tab[i] = i < 24 ? a.e_chrg[i] : (i < 80 ? a.e_pdg[i - 24] : a.e_pv[i - 80]);
    __syncthreads();

    Outer<H, H> oA; oA.zero();
    Outer<16, 8> oC; oC.zero();
    Outer<16, 24> oK; oK.zero();
    ColSum<H> cA; cA.zero();
    ColSum<16> cC; cC.zero();
    ColSum<16> cK; cK.zero();
    float tab_acc[2] = {0.f, 0.f};

    const int tiles = (a.N + NODE_TILE - 1) / NODE_TILE;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const int node0 = tile * NODE_TILE, rows = min(NODE_TILE, a.N - node0);
        float xc[8], du[H], hact[16], ecat[16];
        int ic = 0, ip = 0, iv = 0;
        if (t < rows) {
            const size_t n = node0 + t;
            const float *row = a.x_cont + n * a.x_cont_stride;
#pragma unroll
            for (int k = 0; k < 8; ++k) xc[k] = k < a.CD ? row[k] : 0.f;
            long long c0, c1, c2;
            if (a.x_cat_i64) { const long long *cr = a.x_cat_i64 + n * 3; c0 = cr[0]; c1 = cr[1]; c2 = cr[2]; }
            else { const float *cr = a.x_cat_f32 + n * a.x_cat_stride; c0 = (long long)cr[0]; c1 = (long long)cr[1]; c2 = (long long)cr[2]; }
            long long lc = c1 + 1, lp = pdg_remap(c0), lv = c2;   // clamp exactly like the forward
            ic = (int)(lc < 0 ? 0 : (lc > 2 ? 2 : lc));
            ip = (int)(lp < 0 ? 0 : (lp > 6 ? 6 : lp));
            iv = (int)(lv < 0 ? 0 : (lv > 7 ? 7 : lv));
            const int combo = (ic * 7 + ip) * 8 + iv;
            float u[H], dz[H], a0[H], h0[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) { ecat[k] = s_lut[combo * 16 + k]; u[k] = ecat[k]; }
            dense_smem<8, 16>(s_cont_w, s_cont_b, xc, h0);
#pragma unroll
            for (int k = 0; k < 16; ++k) { hact[k] = elu(h0[k]); u[16 + k] = hact[k]; }
            load_row32(a.enc + n * H, a0);
            load_row32(a.DA + n * H, dz);
#pragma unroll
            for (int k = 0; k < H; ++k) dz[k] *= elu_grad_from_out(a0[k]);
            store_row<H>(tA + t * GLD, dz);
            store_row<H>(tB + t * GLD, u);
            dense_t_smem<H, H>(s_enc_w, dz, du);
        }
        __syncthreads();
        oA.accumulate(tA, GLD, tB, GLD, rows);
        cA.accumulate(tA, GLD, rows);
        __syncthreads();
        float dpk[16];
        if (t < rows) {
            float dh0[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                dh0[k] = du[16 + k] * elu_grad_from_out(hact[k]);
                dpk[k] = du[k] * elu_grad_from_out(ecat[k]);
            }
            store_row<16>(tA + t * GLD, dh0);
            store_row<8>(tB + t * GLD, xc);
        }
        __syncthreads();
        oC.accumulate(tA, GLD, tB, GLD, rows);
        cC.accumulate(tA, GLD, rows);
        __syncthreads();
        if (t < rows) {
            float v[24];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                v[k] = s_tab[ic * 8 + k]; v[8 + k] = s_tab[24 + ip * 8 + k]; v[16 + k] = s_tab[80 + iv * 8 + k];
            }
            store_row<16>(tA + t * GLD, dpk);
            store_row<24>(tB + t * GLD, v);
        }
        __syncthreads();
        oK.accumulate(tA, GLD, tB, GLD, rows);
        cK.accumulate(tA, GLD, rows);
        __syncthreads();
        if (t < rows) {
            float dv[24];
            dense_t_smem<24, 16>(s_cat_w, dpk, dv);
            store_row<24>(tB + t * GLD, dv);
            s_idx[t] = ic | (ip << 8) | (iv << 16);
        }
        __syncthreads();
        // embedding-table gradients: accumulator (row r of the stacked 18-row table, column k)
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int id = t + q * NODE_TILE;
            if (id < 144) {
                const int r = id >> 3, k = id & 7;
                const int shift = r < 3 ? 0 : (r < 10 ? 8 : 16);
                const int want = r < 3 ? r : (r < 10 ? r - 3 : r - 10);
                const int colo = (r < 3 ? 0 : (r < 10 ? 8 : 16)) + k;
                float acc = tab_acc[q];
                for (int rr = 0; rr < rows; ++rr)
                    if (((s_idx[rr] >> shift) & 0xff) == want) acc += tB[rr * GLD + colo];
                tab_acc[q] = acc;
            }
        }
        __syncthreads();
    }
    float *dst = a.gpart + (size_t)blockIdx.x * GP_STRIDE;
    oA.flush(scratch, dst);
    cA.flush(scratch, dst + 1024, H);
    oC.flush(scratch, dst + 1056);
    cC.flush(scratch, dst + 1184, 16);
    oK.flush(scratch, dst + 1200);
    cK.flush(scratch, dst + 1584, 16);
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        const int id = t + q * NODE_TILE;
        if (id < 144) dst[1600 + id] = tab_acc[q];
    }
}

constexpr size_t CONV_BWD_SMEM = (size_t)(2 * H * H + 3 * NODE_TILE * GLD) * 4;
constexpr size_t ENC_BWD_SMEM = (size_t)(H * H + 128 + 16 + 384 + LUT_ROWS * 16 + 144 + 2 * NODE_TILE * GLD + 1024 + NODE_TILE) * 4;

int bn_backward(const float *G, const float *raw, const float *stat, int N, int training, float *dgamma, float *dbeta,
                float *DC, const BwdWs &w, cudaStream_t st) {
    const size_t n4 = (size_t)N * (H / 4);
    const int grid = persistent_grid((long long)((n4 + BNB_THREADS - 1) / BNB_THREADS), 2);
    k_bn_bwd_reduce<<<grid, BNB_THREADS, 0, st>>>((const float4 *)G, (const float4 *)raw, stat, n4, w.bnpart);
    GM_LAUNCH_CHECK();
    const int grid2 = persistent_grid((long long)((n4 + BNB_THREADS - 1) / BNB_THREADS), 8);
    k_bn_bwd_apply<<<grid2, BNB_THREADS, 0, st>>>((const float4 *)G, (const float4 *)raw, stat, w.bnpart, grid, N,
                                                  training, dgamma, dbeta, (float4 *)DC, n4);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

}  // namespace

int check_params(const GmParams *p);

}  // namespace gm

using namespace gm;

extern "C" {

size_t gm_net_bwd_ws_bytes(int N, int conv_depth) {
    (void)conv_depth;
    return bwd_ws_layout(N < 0 ? 0 : N, nullptr, nullptr);
}

int gm_net_backward(const GmParams *p, const GmParams *grads, const float *x_cont, int x_cont_stride,
                    const int64_t *x_cat_i64, const float *x_cat_f32, int x_cat_stride, const int *rowptr,
                    const int *col, const int *rowptr_t, const int *col_t, const int *use_transposed_flag, int N,
                    int training, int apply_sigmoid, int aggr, float eps, const void *fwd_ws, size_t fwd_ws_bytes, void *bwd_ws, size_t bwd_ws_bytes,
                    const float *out, const float *grad_out, void *stream) {
    (void)eps;
    GM_TRY(check_params(p));
    if (!grads || N < 0 || !fwd_ws || !bwd_ws) return GM_ERR_INVALID_ARGUMENT;
    if (aggr != GM_AGGR_MAX) return GM_ERR_UNSUPPORTED;
    if (N > 0 && (!x_cont || (!x_cat_i64 && !x_cat_f32) || !rowptr_t || !col_t || !out || !grad_out))
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
    const int ngrid = persistent_grid(tiles, 2);

    // ---- head
    k_head_bwd<<<ngrid, NODE_TILE, 0, st>>>(f.emb[depth], out, grad_out, N, p->out1_w, p->out1_b, p->out2_w,
                                            apply_sigmoid, w.G, w.gpart);
    GM_LAUNCH_CHECK();
    {
        Segs s{};
        s.n = 4;
        s.s[0] = Seg{grads->out1_w, H, 0, 16, H, H};
        s.s[1] = Seg{grads->out1_b, 16, 512, 1, 16, 16};
        s.s[2] = Seg{grads->out2_w, 16, 528, 1, 16, 16};
        s.s[3] = Seg{grads->out2_b, 1, 544, 1, 1, 1};
        GM_TRY(launch_reduce(w.gpart, ngrid, s, st));
    }
    // ---- EdgeConv layers, last to first
    for (int l = depth - 1; l >= 0; --l) {
        const float *stat = f.bnstat + (size_t)(l + 1) * BNSTAT_STRIDE;
        GM_TRY(bn_backward(w.G, f.conv[l], stat, N, training, grads->conv_bn[l].weight, grads->conv_bn[l].bias, w.DC, w, st));
        {
            TimedRegion tr(GM_TIMER_SCATTER_S_BWD, st);
            k_scatter_s<<<persistent_grid(((long long)N * 32 + 255) / 256, 8), 256, 0, st>>>(
                w.DC, f.arg[l], rowptr, col, rowptr_t, col_t, use_transposed_flag, N, w.S);
            GM_LAUNCH_CHECK();
        }
        GM_CUDA_TRY(cudaFuncSetAttribute(k_conv_bwd_node, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CONV_BWD_SMEM));
        {
            TimedRegion tr(GM_TIMER_CONV_NODE_BWD, st);
            k_conv_bwd_node<<<ngrid, NODE_TILE, CONV_BWD_SMEM, st>>>(f.emb[l], w.DC, w.S, N, p->conv_w[l], w.G, w.gpart);
            GM_LAUNCH_CHECK();
        }
        Segs s{};
        s.n = 3;
        s.s[0] = Seg{grads->conv_w[l], CONV_IN, 0, H, H, H};
        s.s[1] = Seg{grads->conv_w[l] ? grads->conv_w[l] + H : nullptr, CONV_IN, 1024, H, H, H};
        s.s[2] = Seg{grads->conv_b[l], H, 2048, 1, H, H};
        GM_TRY(launch_reduce(w.gpart, ngrid, s, st));
    }
    // ---- bn_all + encoder
    GM_TRY(bn_backward(w.G, f.enc, f.bnstat, N, training, grads->bn_all.weight, grads->bn_all.bias, w.DC, w, st));
    EncBwdArgs ea{};
    ea.x_cont = x_cont; ea.x_cont_stride = x_cont_stride; ea.x_cat_i64 = (const long long *)x_cat_i64;
    ea.x_cat_f32 = x_cat_f32; ea.x_cat_stride = x_cat_stride; ea.N = N; ea.CD = p->continuous_dim;
    ea.cont_w = p->cont_w; ea.cont_b = p->cont_b; ea.cat_w = p->cat_w; ea.enc_w = p->enc_w; ea.lut = f.lut;
    ea.e_chrg = p->embed_charge; ea.e_pdg = p->embed_pdgid; ea.e_pv = p->embed_pv;
    ea.enc = f.enc; ea.DA = w.DC; ea.gpart = w.gpart;
    GM_CUDA_TRY(cudaFuncSetAttribute(k_encode_bwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ENC_BWD_SMEM));
    {
        TimedRegion tr(GM_TIMER_ENCODE_BWD, st);
        k_encode_bwd<<<ngrid, NODE_TILE, ENC_BWD_SMEM, st>>>(ea);
        GM_LAUNCH_CHECK();
    }
    {
        const int CD = p->continuous_dim;
        Segs s{};
        s.n = 9;
        s.s[0] = Seg{grads->enc_w, H, 0, H, H, H};
        s.s[1] = Seg{grads->enc_b, H, 1024, 1, H, H};
        s.s[2] = Seg{grads->cont_w, CD, 1056, 16, 8, CD};
        s.s[3] = Seg{grads->cont_b, 16, 1184, 1, 16, 16};
        s.s[4] = Seg{grads->cat_w, 24, 1200, 16, 24, 24};
        s.s[5] = Seg{grads->cat_b, 16, 1584, 1, 16, 16};
        s.s[6] = Seg{grads->embed_charge, 8, 1600, 3, 8, 8};
        s.s[7] = Seg{grads->embed_pdgid, 8, 1624, 7, 8, 8};
        s.s[8] = Seg{grads->embed_pv, 8, 1680, 8, 8, 8};
        GM_TRY(launch_reduce(w.gpart, ngrid, s, st));
    }
    return GM_OK;
}

}  // extern "C"
