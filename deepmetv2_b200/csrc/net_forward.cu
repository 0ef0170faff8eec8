// Forward of the GraphMETNetwork hot path (model/graph_met_network.py:47-69, model/net.py:45-47):
//   k_build_lut   : the 168-entry categorical table ELU(Linear_24->16(cat[E_chrg,E_pdg,E_pv]))
//   k_encode2     : fused embeddings + continuous encoder + encode_all (+ BatchNorm moments)   [a3,a4]
//   k_node_prep   : emb_l = (resid +) BN(raw) ; Y = W2 emb_l ; P = (W1-W2) emb_l + b            [K4 + a5, node side]
//   k_edge_aggr   : out_i = P_i + max_j Y_j over the CSR row (segmented, atomics-free) + BN moments [a5, edge side]
//   k_edgeconv_simt: fp32 EdgeConv with the per-edge contraction W2 (x_j - x_i) spelled out (cross-check variant)
//   k_head        : emb_L = resid + BN(raw) ; Linear_32->16 + ELU + Linear_16->1 (+ sigmoid)    [a6]
// BatchNorm statistics (train mode) are combined by the LAST CTA of the producing kernel in a fixed order
// (bn_finalize_cta): no separate launch, no CTA ever waits on another, bit-deterministic.
#include "geo_layout.cuh"
#include "net_layout.cuh"

namespace gm {

int edgeconv_tc_forward(const float *x, const int *rowptr, const int *col, int N, const float *W, const float *b,
                        int aggr, int precision, float *out, int *arg, float *partials, int *n_partials,
                        cudaStream_t st);

namespace {

// ------------------------------------------------------------------------------------------ LUT
__global__ void __launch_bounds__(256) k_build_lut(GmParams p, float *__restrict__ lut, int *__restrict__ counters,
                                                   float *__restrict__ bnstat, int training, float eps) {
    pdl_sync();
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx < NUM_COUNTERS) counters[idx] = 0;
    if (!training && blockIdx.x == 0) {      // eval mode: BatchNorm is a fixed affine of the running statistics
        for (int k = threadIdx.x; k < (p.conv_depth + 1) * H; k += blockDim.x) {
            const int l = k >> 5, c = k & 31;
            const GmBatchNorm &bn = l == 0 ? p.bn_all : p.conv_bn[l - 1];
            const float invstd = 1.0f / sqrtf(bn.running_var[c] + eps);
            float *stat = bnstat + (size_t)l * BNSTAT_STRIDE;
            stat[c] = bn.running_mean[c];
            stat[H + c] = bn.weight[c] * invstd;
            stat[2 * H + c] = invstd;
            stat[3 * H + c] = bn.bias[c];
        }
    }
    if (idx >= LUT_ROWS * 16) return;
    const int o = idx & 15, combo = idx >> 4;
    const int pv = combo & 7, pdg = (combo >> 3) % 7, ch = (combo >> 3) / 7;
    // concat order charge, pdg, pv (model/graph_met_network.py:57)
    float acc = p.cat_b[o];
    const float *w = p.cat_w + o * 24;
    for (int k = 0; k < 8; ++k) acc = fmaf(w[k], p.embed_charge[ch * 8 + k], acc);
    for (int k = 0; k < 8; ++k) acc = fmaf(w[8 + k], p.embed_pdgid[pdg * 8 + k], acc);
    for (int k = 0; k < 8; ++k) acc = fmaf(w[16 + k], p.embed_pv[pv * 8 + k], acc);
    lut[idx] = elu(acc);
}

__device__ __forceinline__ void store_tile(const float *tile, float *__restrict__ out, int node0, int rows) {
    // coalesced float4 stores of a [rows][32] tile
    for (int idx = threadIdx.x; idx < rows * (H / 4); idx += blockDim.x) {
        int r = idx >> 3, c4 = (idx & 7) * 4;
        const float *s = tile + r * TILE_LD + c4;
        reinterpret_cast<float4 *>(out + (size_t)(node0 + r) * H)[idx & 7] = make_float4(s[0], s[1], s[2], s[3]);
    }
}

// ------------------------------------------------------------------------------------------ encoder
struct EncodeArgs {
    const float *x_cont; int x_cont_stride;
    const long long *x_cat_i64;
    const float *x_cat_f32; int x_cat_stride;
    const int *perm;          // internal position -> input row (NULL = identity)
    int N, CD, flat;          // flat: x_cont/x_cat_f32 are columns of one [N, stride] matrix (stride <= 16), no perm
    const float *cont_w, *cont_b, *enc_w, *enc_b, *lut;
    float *enc_out;           // [N,32]
    float *partials;          // [grid][3][32] or NULL (eval mode)
    int *counter;             // last-CTA ticket (train mode)
    GmBatchNorm bn; float momentum, eps; float *stat;
    int *err_flags;
};

__device__ __forceinline__ int cat_combo(long long pdg, long long chg, long long pv, int *err) {
    long long ic = chg + 1, ip = pdg_remap(pdg), iv = pv;
    if (ic < 0 || ic > 2 || ip < 0 || ip > 6 || iv < 0 || iv > 7) {
        *err = 1;
        ic = ic < 0 ? 0 : (ic > 2 ? 2 : ic);
        ip = ip < 0 ? 0 : (ip > 6 ? 6 : ip);
        iv = iv < 0 ? 0 : (iv > 7 ? 7 : iv);
    }
    return (int)((ic * 7 + ip) * 8 + iv);
}

// Encoder on the register-blocked tile product: 256 threads per 128-node tile.  Threads 0..127 build the node's
// 32-vector u = [LUT row | ELU(W_c x_cont + b)] (thread per node: 256 flop), then ALL threads run the 32x32 layer
// (2/3 of the encoder's flops) as tile_gemm (4 nodes x 4 outputs per thread, K-major weights), apply ELU and store
// whole 128-byte rows from registers.  BatchNorm moments: results go back into the tile, 8 slices of 16 rows.
constexpr int ENC_THREADS = 256;
__global__ void __launch_bounds__(ENC_THREADS, 3) k_encode2(EncodeArgs a) {
    __shared__ __align__(16) float s_cont_w[16 * 8];
    __shared__ float s_cont_b[16];
    __shared__ __align__(16) float s_encT[H * H];                // enc_w transposed: [k][o]
    __shared__ __align__(16) float s_enc_b[H];
    __shared__ __align__(16) float s_lut[LUT_ROWS * 16];
    __shared__ __align__(16) float s_x[NODE_TILE * 16];
    __shared__ float s_tile[NODE_TILE * TILE_LD];
    __shared__ float s_red[ENC_THREADS / 32][3][H];              // 8 slices for the moments
    static_assert(NODE_TILE * TILE_LD >= (ENC_THREADS / 8) * 3 * H, "the finalize scratch aliases the tile");

    const int t = threadIdx.x, ng = t >> 3, og = t & 7;
    for (int i = t; i < 16 * 8; i += ENC_THREADS) {
        int o = i >> 3, k = i & 7;
        s_cont_w[i] = k < a.CD ? a.cont_w[o * a.CD + k] : 0.f;
    }
    if (t < 16) s_cont_b[t] = a.cont_b[t];
    for (int i = t; i < H * H; i += ENC_THREADS) s_encT[i] = a.enc_w[(i & 31) * H + (i >> 5)];
    if (t < H) s_enc_b[t] = a.enc_b[t];
    pdl_sync();                                  // weights above: parameters; the LUT is the previous kernel's output
    for (int i = t; i < LUT_ROWS * 16; i += ENC_THREADS) s_lut[i] = a.lut[i];
    __syncthreads();
    const float4 b4 = reinterpret_cast<const float4 *>(s_enc_b)[og];

    Moments run{0.f, 0.f, 0.f};
    int err = 0;
    const int tiles = (a.N + NODE_TILE - 1) / NODE_TILE;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const int node0 = tile * NODE_TILE;
        const int rows = min(NODE_TILE, a.N - node0);
        float xc[8];
        long long c0 = 0, c1 = 0, c2 = 0;
        if (a.flat) {
            const int S = a.x_cont_stride;
            const float *src = a.x_cont + (size_t)node0 * S;
            for (int i = t; i < rows * S; i += ENC_THREADS) s_x[i] = src[i];
            __syncthreads();
            if (t < rows) {
                const float *row = s_x + t * S;
                const int cat_off = (int)(a.x_cat_f32 - a.x_cont);
#pragma unroll
                for (int k = 0; k < 8; ++k) xc[k] = k < a.CD ? row[k] : 0.f;
                c0 = (long long)row[cat_off]; c1 = (long long)row[cat_off + 1]; c2 = (long long)row[cat_off + 2];
            }
        } else if (t < rows) {
            const size_t n = a.perm ? (size_t)a.perm[node0 + t] : (size_t)(node0 + t);
            const float *row = a.x_cont + n * a.x_cont_stride;
#pragma unroll
            for (int k = 0; k < 8; ++k) xc[k] = k < a.CD ? row[k] : 0.f;
            if (a.x_cat_i64) {
                const long long *cr = a.x_cat_i64 + n * 3;
                c0 = cr[0]; c1 = cr[1]; c2 = cr[2];
            } else {
                const float *cr = a.x_cat_f32 + n * a.x_cat_stride;
                c0 = (long long)cr[0]; c1 = (long long)cr[1]; c2 = (long long)cr[2];
            }
        }
        if (t < rows) {                                           // u = cat[emb_cat, emb_cont] (graph_met_network.py:58)
            float *u = s_tile + t * TILE_LD;
            const int combo = cat_combo(c0, c1, c2, &err);
            const float4 *l4 = reinterpret_cast<const float4 *>(s_lut + combo * 16);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float4 v = l4[k];
                u[4 * k] = v.x; u[4 * k + 1] = v.y; u[4 * k + 2] = v.z; u[4 * k + 3] = v.w;
            }
            float h0[16];
            dense_smem<8, 16>(s_cont_w, s_cont_b, xc, h0);
#pragma unroll
            for (int k = 0; k < 16; ++k) u[16 + k] = elu(h0[k]);
        } else if (t < NODE_TILE) {                               // rows past the end: keep the product finite
            float *u = s_tile + t * TILE_LD;
#pragma unroll
            for (int k = 0; k < H; ++k) u[k] = 0.f;
        }
        __syncthreads();
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { acc[i][0] = b4.x; acc[i][1] = b4.y; acc[i][2] = b4.z; acc[i][3] = b4.w; }
        tile_gemm<H, H, TILE_LD>(s_tile, s_encT, ng, og, acc);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = elu(acc[i][j]);
            if (ng + 32 * i < rows)
                reinterpret_cast<float4 *>(a.enc_out)[(size_t)(node0 + ng + 32 * i) * (H / 4) + og] =
                    make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        }
        __syncthreads();                                          // every thread is done reading u
        if (a.partials) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float *d = s_tile + (ng + 32 * i) * TILE_LD + 4 * og;
                d[0] = acc[i][0]; d[1] = acc[i][1]; d[2] = acc[i][2]; d[3] = acc[i][3];
            }
            __syncthreads();
            const int c = t & 31, q = t >> 5;                     // slice q: rows 16q .. 16q+15
            const int r0 = q * 16, r1 = min(rows, r0 + 16);
            float n = (float)max(r1 - r0, 0), mean = 0.f, m2 = 0.f;
            if (r1 > r0) {
                float sum = 0.f;
                for (int r = r0; r < r1; ++r) sum += s_tile[r * TILE_LD + c];
                mean = sum / n;
                for (int r = r0; r < r1; ++r) { float dlt = s_tile[r * TILE_LD + c] - mean; m2 = fmaf(dlt, dlt, m2); }
            }
            s_red[q][0][c] = n; s_red[q][1][c] = mean; s_red[q][2][c] = m2;
            __syncthreads();
            if (t < 32) {
#pragma unroll
                for (int gq = 0; gq < ENC_THREADS / 32; ++gq)
                    run = moments_combine(run, Moments{s_red[gq][0][c], s_red[gq][1][c], s_red[gq][2][c]});
            }
            __syncthreads();
        }
    }
    if (err) atomicOr(a.err_flags, GM_FLAG_CAT_OUT_OF_RANGE);
    if (a.partials) {
        if (t < 32) {
            float *dst = a.partials + (size_t)blockIdx.x * MOM_STRIDE;
            dst[t] = run.n; dst[H + t] = run.mean; dst[2 * H + t] = run.m2;
        }
        if (last_cta_done(a.counter))
            bn_finalize_cta(a.partials, gridDim.x, a.bn, a.momentum, a.eps, a.stat,
                            reinterpret_cast<float(*)[3][H]>(s_tile));
    }
}

// ------------------------------------------------------------------------------------------ BatchNorm (TC path only)
// Stand-alone finalize for producers that do not carry the last-CTA epilogue (the tcgen05 EdgeConv kernel).
__global__ void __launch_bounds__(256) k_bn_finalize(const float *__restrict__ partials, int n_partials,
                                                     GmBatchNorm bn, float momentum, float eps,
                                                     float *__restrict__ stat) {
    __shared__ float red[32][3][H];
    bn_finalize_cta(partials, n_partials, bn, momentum, eps, stat, red);
}

// ------------------------------------------------------------------------------------------ EdgeConv, node side
// The single-Linear edge MLP of the reference (model/graph_met_network.py:34-38) splits exactly:
//     W [x_i ; x_j - x_i] + b  =  (W1 - W2) x_i + b  +  W2 x_j   =  P_i + Y_j ,
// so max_j / sum_j only ever touch Y_j.  This kernel materialises, in ONE pass over the nodes,
//     emb = (resid +) BN(raw)      the layer input (previous BatchNorm + residual folded in, K4)
//     Y   = W2 emb                 what the edge side gathers
//     P   = (W1 - W2) emb + b      the per-target term
// Thread per node, weights broadcast from shared memory, coalesced tile loads/stores.
template <bool WRITE_YP>
__global__ void __launch_bounds__(NODE_TILE)
k_node_prep(const float *__restrict__ raw, const float *__restrict__ resid, const float *__restrict__ stat,
            const float *__restrict__ W, const float *__restrict__ bias, int N, float *__restrict__ emb_out,
            float *__restrict__ Y, float *__restrict__ P) {
    __shared__ __align__(16) float s_wd[H * H];     // W1 - W2, [c][k]
    __shared__ __align__(16) float s_w2[H * H];
    __shared__ float s_b[H], s_zero[H];
    __shared__ float s_stat[BNSTAT_STRIDE];
    __shared__ float s_tile[NODE_TILE * TILE_LD];
    const int t = threadIdx.x;
    if (WRITE_YP) {
        for (int i = t; i < H * H; i += NODE_TILE) {
            const int c = i >> 5, k = i & 31;
            const float w1 = W[c * CONV_IN + k], w2 = W[c * CONV_IN + H + k];
            s_wd[i] = w1 - w2; s_w2[i] = w2;
        }
        if (t < H) { s_b[t] = bias[t]; s_zero[t] = 0.f; }
    }
    if (stat && t < BNSTAT_STRIDE) s_stat[t] = stat[t];
    __syncthreads();
    const int tiles = (N + NODE_TILE - 1) / NODE_TILE;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const int node0 = tile * NODE_TILE, rows = min(NODE_TILE, N - node0);
        // coalesced load with the BatchNorm affine + residual applied on the fly; emb is stored right away
        for (int idx = t; idx < rows * (H / 4); idx += NODE_TILE) {
            const int r = idx >> 3, c4 = (idx & 7) * 4;
            const size_t g = (size_t)(node0 + r) * (H / 4) + (idx & 7);
            float4 v = reinterpret_cast<const float4 *>(raw)[g];
            if (stat) {
                v.x = fmaf(v.x - s_stat[c4 + 0], s_stat[H + c4 + 0], s_stat[3 * H + c4 + 0]);
                v.y = fmaf(v.y - s_stat[c4 + 1], s_stat[H + c4 + 1], s_stat[3 * H + c4 + 1]);
                v.z = fmaf(v.z - s_stat[c4 + 2], s_stat[H + c4 + 2], s_stat[3 * H + c4 + 2]);
                v.w = fmaf(v.w - s_stat[c4 + 3], s_stat[H + c4 + 3], s_stat[3 * H + c4 + 3]);
            }
            if (resid) {
                const float4 q = reinterpret_cast<const float4 *>(resid)[g];
                v.x += q.x; v.y += q.y; v.z += q.z; v.w += q.w;
            }
            if (emb_out) reinterpret_cast<float4 *>(emb_out)[g] = v;
            float *d = s_tile + r * TILE_LD + c4;
            d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
        }
        __syncthreads();
        if (WRITE_YP) {
            float x[H], y[H];
            if (t < rows) {
#pragma unroll
                for (int k = 0; k < H; ++k) x[k] = s_tile[t * TILE_LD + k];
                dense_smem<H, H>(s_w2, s_zero, x, y);
#pragma unroll
                for (int k = 0; k < H; ++k) s_tile[t * TILE_LD + k] = y[k];
            }
            __syncthreads();
            store_tile(s_tile, Y, node0, rows);
            __syncthreads();
            if (t < rows) {
                dense_smem<H, H>(s_wd, s_b, x, y);
#pragma unroll
                for (int k = 0; k < H; ++k) s_tile[t * TILE_LD + k] = y[k];
            }
            __syncthreads();
            store_tile(s_tile, P, node0, rows);
        }
        __syncthreads();
    }
}

// Same layer-input pass on the register-blocked tile product (common.cuh: tile_gemm): 256 threads per 128-node tile,
// thread = 4 nodes x 4 outputs, weights K-major in shared memory, results stored straight from registers as whole
// 128-byte rows.  ~64 registers instead of ~100 and 16 independent FMA chains per thread instead of one: the
// thread-per-node version above ran at 23% of the FMA pipe with 16 warps per SM (profiles/r01c_ncu_node_kernels.txt).
constexpr int NP_THREADS = 256;
__global__ void __launch_bounds__(NP_THREADS, 4)
k_node_prep2(const float *__restrict__ raw, const float *__restrict__ resid, const float *__restrict__ stat,
             const float *__restrict__ W, const float *__restrict__ bias, int N, float *__restrict__ emb_out,
             float *__restrict__ Y, float *__restrict__ P, int direct = 0) {
    // direct != 0 (per-edge tcgen05 path, geo_tc.cu): P = W1 x + b and no Y (pass Y == NULL)
    __shared__ __align__(16) float s_wdT[H * H];    // (W1 - W2)^T : [k][c]
    __shared__ __align__(16) float s_w2T[H * H];    // W2^T
    __shared__ __align__(16) float s_b[H];
    __shared__ float s_stat[BNSTAT_STRIDE];
    __shared__ float s_x[NODE_TILE * TILE_LD];
    const int t = threadIdx.x, ng = t >> 3, og = t & 7;
    for (int i = t; i < H * H; i += NP_THREADS) {
        const int k = i >> 5, c = i & 31;
        const float w1 = W[c * CONV_IN + k], w2 = W[c * CONV_IN + H + k];
        s_wdT[i] = direct ? w1 : w1 - w2; s_w2T[i] = w2;
    }
    if (t < H) s_b[t] = bias[t];
    pdl_sync();                                  // weights above: parameters; stat is the previous kernel's output
    if (stat && t < BNSTAT_STRIDE) s_stat[t] = stat[t];
    __syncthreads();
    const float4 b4 = reinterpret_cast<const float4 *>(s_b)[og];
    const int tiles = (N + NODE_TILE - 1) / NODE_TILE;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const int node0 = tile * NODE_TILE, rows = min(NODE_TILE, N - node0);
        // all loads of the tile first (8 x 16 bytes per thread in flight), then the arithmetic: ONE memory round trip per
        // tile -- the rolled loop issued a load pair per trip and waited for it (four round trips)
        constexpr int NP_IT = NODE_TILE * (H / 4) / NP_THREADS;
        float4 rv[NP_IT], qv[NP_IT];
#pragma unroll
        for (int it = 0; it < NP_IT; ++it) {
            const int idx = it * NP_THREADS + t, r = idx >> 3;
            const size_t g = (size_t)(node0 + r) * (H / 4) + (idx & 7);
            rv[it] = qv[it] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (r < rows) {
                rv[it] = reinterpret_cast<const float4 *>(raw)[g];
                if (resid) qv[it] = reinterpret_cast<const float4 *>(resid)[g];
            }
        }
#pragma unroll
        for (int it = 0; it < NP_IT; ++it) {
            const int idx = it * NP_THREADS + t, r = idx >> 3, c4 = (idx & 7) * 4;
            if (r >= rows) continue;
            const size_t g = (size_t)(node0 + r) * (H / 4) + (idx & 7);
            float4 v = rv[it];
            if (stat) {
                v.x = fmaf(v.x - s_stat[c4 + 0], s_stat[H + c4 + 0], s_stat[3 * H + c4 + 0]);
                v.y = fmaf(v.y - s_stat[c4 + 1], s_stat[H + c4 + 1], s_stat[3 * H + c4 + 1]);
                v.z = fmaf(v.z - s_stat[c4 + 2], s_stat[H + c4 + 2], s_stat[3 * H + c4 + 2]);
                v.w = fmaf(v.w - s_stat[c4 + 3], s_stat[H + c4 + 3], s_stat[3 * H + c4 + 3]);
            }
            if (resid) {
                const float4 q = qv[it];
                v.x += q.x; v.y += q.y; v.z += q.z; v.w += q.w;
            }
            if (emb_out) reinterpret_cast<float4 *>(emb_out)[g] = v;
            float *d = s_x + r * TILE_LD + c4;
            d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
        }
        __syncthreads();
        float acc[4][4];
        if (Y) {
#pragma unroll
            for (int i = 0; i < 4; ++i) { acc[i][0] = 0.f; acc[i][1] = 0.f; acc[i][2] = 0.f; acc[i][3] = 0.f; }
            tile_gemm<H, H, TILE_LD>(s_x, s_w2T, ng, og, acc);
#pragma unroll
            for (int i = 0; i < 4; ++i)
                if (ng + 32 * i < rows)
                    reinterpret_cast<float4 *>(Y)[(size_t)(node0 + ng + 32 * i) * (H / 4) + og] =
                        make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) { acc[i][0] = b4.x; acc[i][1] = b4.y; acc[i][2] = b4.z; acc[i][3] = b4.w; }
        tile_gemm<H, H, TILE_LD>(s_x, s_wdT, ng, og, acc);
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (ng + 32 * i < rows)
                reinterpret_cast<float4 *>(P)[(size_t)(node0 + ng + 32 * i) * (H / 4) + og] =
                    make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------ EdgeConv, edge side
// out_i = P_i + max_{j in N(i)} Y_j   (add: deg*P_i + sum_j Y_j ; mean: P_i + mean_j Y_j)
// Warp per target, HALF-WARP per source row: a 128-byte Y row is read by 16 lanes as float2, so one LDG.64
// of the warp moves two source rows and every lane keeps the running max / arg-max (or sum) of 2 channels
// over the edges of its half; up to 16 loads (32 edges) are in flight per lane and the next target's row
// bounds / sources are prefetched while the current one is reduced.  The per-target epilogue is one
// xor-16 exchange.  Rows are read in CSR order (ascending source) with a strict '>' and the halves are
// merged with "lowest source index wins", so ties are broken exactly like a sequential scan: a segmented,
// fixed-order, atomics-free reduction.  BatchNorm moments of the output ride along as shifted sums
// (shift = the warp's first output, so nothing cancels) and the last CTA finalises them.
// [profiles/r01_ncu_edge_aggr.txt: the float4 / 8-lanes-per-row variant spent 48% of its instructions in
//  the per-target epilogue at the DY/TT degree of ~27; this layout cuts the epilogue to one exchange.]
constexpr int EA_WARPS = 8;


template <int AGGR, bool ARG>
__device__ __forceinline__ void aggr_consume(float2 &acc, int2 &barg, const float2 (&v)[8], const int (&j)[8]) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
        if (AGGR == GM_AGGR_MAX && ARG) {
            if (v[u].x > acc.x) { acc.x = v[u].x; barg.x = j[u]; }
            if (v[u].y > acc.y) { acc.y = v[u].y; barg.y = j[u]; }
        } else if (AGGR == GM_AGGR_MAX) {
            acc.x = fmaxf(acc.x, v[u].x); acc.y = fmaxf(acc.y, v[u].y);
        } else {
            acc.x += v[u].x; acc.y += v[u].y;
        }
    }
}

template <int AGGR, bool ARG>
__global__ void __launch_bounds__(EA_WARPS * 32)
k_edge_aggr(const float *__restrict__ Y, const float *__restrict__ P, const int *__restrict__ rowptr,
            const int *__restrict__ col, int N, float *__restrict__ out, int *__restrict__ arg, AggrTail tail) {
    __shared__ float s_red[EA_WARPS * 4][3][H];   // per-warp moments (EA_WARPS rows), then the finalize's lane groups
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane & 15, half = lane >> 4;
    const float2 *__restrict__ Y2 = reinterpret_cast<const float2 *>(Y);
    const float NEG = -INFINITY;
    const float2 IDENT = AGGR == GM_AGGR_MAX ? make_float2(NEG, NEG) : make_float2(0.f, 0.f);
    // shifted sums for the BatchNorm moments of this warp's outputs (lanes of half 0; 2 channels each)
    float cnt = 0.f;
    float2 shift = make_float2(0.f, 0.f), s1 = shift, s2 = shift;
    const int stride = gridDim.x * EA_WARPS;
    int i = blockIdx.x * EA_WARPS + warp;
    // software pipeline over targets: row bounds and the first 32 sources of the NEXT target are fetched
    // while the current one is reduced
    int beg = 0, end = 0, jl = -1;
    if (i < N) {
        beg = __ldg(rowptr + i); end = __ldg(rowptr + i + 1);
        jl = lane < end - beg ? __ldg(col + beg + lane) : -1;
    }
    while (i < N) {
        const int inext = i + stride;
        int nbeg = 0, nend = 0, jnext = -1;
        if (inext < N) { nbeg = __ldg(rowptr + inext); nend = __ldg(rowptr + inext + 1); }
        float2 p = make_float2(0.f, 0.f);
        if (half == 0) p = __ldg(reinterpret_cast<const float2 *>(P) + (size_t)i * (H / 2) + sub);
        float2 acc = IDENT;
        int2 barg = make_int2(-1, -1);
        for (int e0 = beg; e0 < end; e0 += 32) {
            const int nb = min(32, end - e0);
            if (e0 > beg) jl = lane < nb ? __ldg(col + e0 + lane) : -1;
            // up to 16 row loads in flight per lane (edge 2u + half of this 32-edge chunk)
            int ja[8], jb[8];
            float2 va[8], vb[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) ja[u] = __shfl_sync(0xffffffffu, jl, 2 * u + half);
#pragma unroll
            for (int u = 0; u < 8; ++u) va[u] = (2 * u + half < nb) ? __ldg(Y2 + (size_t)ja[u] * (H / 2) + sub) : IDENT;
            if (nb > 16) {
#pragma unroll
                for (int u = 0; u < 8; ++u) jb[u] = __shfl_sync(0xffffffffu, jl, 16 + 2 * u + half);
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    vb[u] = (16 + 2 * u + half < nb) ? __ldg(Y2 + (size_t)jb[u] * (H / 2) + sub) : IDENT;
            }
            if (e0 == beg && inext < N) jnext = lane < nend - nbeg ? __ldg(col + nbeg + lane) : -1;
            aggr_consume<AGGR, ARG>(acc, barg, va, ja);
            if (nb > 16) aggr_consume<AGGR, ARG>(acc, barg, vb, jb);
        }
        if (end == beg && inext < N) jnext = lane < nend - nbeg ? __ldg(col + nbeg + lane) : -1;
        // fold the two halves
        const float ox = __shfl_xor_sync(0xffffffffu, acc.x, 16), oy = __shfl_xor_sync(0xffffffffu, acc.y, 16);
        if (AGGR == GM_AGGR_MAX) {
            const float mx = fmaxf(acc.x, ox), my = fmaxf(acc.y, oy);
            if (ARG) {   // lowest source index among the halves that hold the maximum
                const int cx = acc.x == mx ? barg.x : 0x7fffffff, cy = acc.y == my ? barg.y : 0x7fffffff;
                barg.x = min(cx, __shfl_xor_sync(0xffffffffu, cx, 16));
                barg.y = min(cy, __shfl_xor_sync(0xffffffffu, cy, 16));
            }
            acc.x = mx; acc.y = my;
        } else {   // (half 0) + (half 1), evaluated identically in both halves
            acc.x = half == 0 ? acc.x + ox : ox + acc.x;
            acc.y = half == 0 ? acc.y + oy : oy + acc.y;
        }
        if (half == 0) {
            const int deg = end - beg;
            float2 o;
            if (deg == 0) o = make_float2(0.f, 0.f);                                  // PyG: empty rows aggregate to 0
            else if (AGGR == GM_AGGR_MAX) o = make_float2(p.x + acc.x, p.y + acc.y);
            else if (AGGR == GM_AGGR_ADD) o = make_float2(fmaf((float)deg, p.x, acc.x), fmaf((float)deg, p.y, acc.y));
            else {
                const float inv = 1.f / (float)deg;
                o = make_float2(fmaf(acc.x, inv, p.x), fmaf(acc.y, inv, p.y));
            }
            reinterpret_cast<float2 *>(out)[(size_t)i * (H / 2) + sub] = o;
            if (ARG) reinterpret_cast<int2 *>(arg)[(size_t)i * (H / 2) + sub] = barg;
            if (tail.partials) {
                if (cnt == 0.f) shift = o;
                cnt += 1.f;
                const float dx = o.x - shift.x, dy = o.y - shift.y;
                s1.x += dx; s1.y += dy;
                s2.x = fmaf(dx, dx, s2.x); s2.y = fmaf(dy, dy, s2.y);
            }
        }
        i = inext; beg = nbeg; end = nend; jl = jnext;
    }
    if (tail.partials) {
        if (half == 0) {   // shifted sums -> (n, mean, M2)
            const float rn = cnt > 0.f ? 1.f / cnt : 0.f;
            s_red[warp][0][2 * sub] = cnt;     s_red[warp][0][2 * sub + 1] = cnt;
            s_red[warp][1][2 * sub] = fmaf(s1.x, rn, shift.x);
            s_red[warp][1][2 * sub + 1] = fmaf(s1.y, rn, shift.y);
            s_red[warp][2][2 * sub] = fmaxf(s2.x - s1.x * s1.x * rn, 0.f);
            s_red[warp][2][2 * sub + 1] = fmaxf(s2.y - s1.y * s1.y * rn, 0.f);
        }
        __syncthreads();
        if (warp == 0) {
            Moments tot{0.f, 0.f, 0.f};
#pragma unroll
            for (int k = 0; k < EA_WARPS; ++k)
                tot = moments_combine(tot, Moments{s_red[k][0][lane], s_red[k][1][lane], s_red[k][2][lane]});
            float *dst = tail.partials + (size_t)blockIdx.x * MOM_STRIDE;
            dst[lane] = tot.n; dst[H + lane] = tot.mean; dst[2 * H + lane] = tot.m2;
        }
        if (last_cta_done(tail.counter))
            bn_finalize_cta(tail.partials, gridDim.x, tail.bn, tail.momentum, tail.eps, tail.stat, s_red);
    }
}

// ------------------------------------------------------------------------------------------ EdgeConv (fp32 CUDA cores)
// out_i[c] = aggr_{j in N(i)} ( W[c,:] . [x_i ; x_j - x_i] + b[c] )   (torch_geometric EdgeConv, aggr = max)
// The x_i half of the contraction is the same for every edge of a target, so it is evaluated once per
// target (p = b + W1 x_i); the per-edge half W2 (x_j - x_i) runs on FFMA2 with 8 edges in flight.
// Warp per target, lane = output channel; sources are read in CSR order so the segmented max / sum is
// a fixed-order, atomics-free reduction; arg-max keeps the lowest source index on ties.
constexpr int EC_WARPS = 8;
constexpr int EC_EDGES = 8;

template <int AGGR, bool ARG>
__global__ void __launch_bounds__(EC_WARPS * 32, 2)
k_edgeconv_simt(const float *__restrict__ x, const int *__restrict__ rowptr, const int *__restrict__ col, int N,
                const float *__restrict__ W, const float *__restrict__ bias, float *__restrict__ out,
                int *__restrict__ arg, float *__restrict__ partials) {
    __shared__ float s_w1t[H * TILE_LD];                              // W1 transposed [k][c]
    __shared__ __align__(16) float s_d[EC_WARPS][H][EC_EDGES];        // (x_j - x_i)[k] for 8 edges
    __shared__ float s_red[EC_WARPS][3][H];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float w2[H];
#pragma unroll
    for (int k = 0; k < H; ++k) w2[k] = W[lane * CONV_IN + H + k];
    for (int i = threadIdx.x; i < H * H; i += blockDim.x) {
        int c = i >> 5, k = i & 31;
        s_w1t[k * TILE_LD + c] = W[c * CONV_IN + k];
    }
    const float b = bias[lane];
    __syncthreads();

    Moments run{0.f, 0.f, 0.f};
    const int total_warps = gridDim.x * EC_WARPS;
    for (int i = blockIdx.x * EC_WARPS + warp; i < N; i += total_warps) {
        const float xi = x[(size_t)i * H + lane];
        float p = b;
#pragma unroll
        for (int k = 0; k < H; ++k) p = fmaf(s_w1t[k * TILE_LD + lane], __shfl_sync(0xffffffffu, xi, k), p);
        const int beg = rowptr[i], end = rowptr[i + 1];
        float best = -INFINITY, sum = 0.f;
        int barg = -1;
        for (int e0 = beg; e0 < end; e0 += EC_EDGES) {
            const int nb = min(EC_EDGES, end - e0);
            const int jl = lane < nb ? col[e0 + lane] : -1;
            float d[EC_EDGES];
#pragma unroll
            for (int e = 0; e < EC_EDGES; ++e) {
                int j = __shfl_sync(0xffffffffu, jl, e);
                d[e] = j >= 0 ? x[(size_t)j * H + lane] - xi : 0.f;
            }
            float4 *dst = reinterpret_cast<float4 *>(&s_d[warp][lane][0]);
            dst[0] = make_float4(d[0], d[1], d[2], d[3]);
            dst[1] = make_float4(d[4], d[5], d[6], d[7]);
            __syncwarp();
            float2 a01 = make_float2(0.f, 0.f), a23 = a01, a45 = a01, a67 = a01;
#pragma unroll
            for (int k = 0; k < H; ++k) {
                const float4 u = *reinterpret_cast<const float4 *>(&s_d[warp][k][0]);
                const float4 v = *reinterpret_cast<const float4 *>(&s_d[warp][k][4]);
                const float2 wk = make_float2(w2[k], w2[k]);
                a01 = __ffma2_rn(wk, make_float2(u.x, u.y), a01);
                a23 = __ffma2_rn(wk, make_float2(u.z, u.w), a23);
                a45 = __ffma2_rn(wk, make_float2(v.x, v.y), a45);
                a67 = __ffma2_rn(wk, make_float2(v.z, v.w), a67);
            }
            __syncwarp();
            const float acc[EC_EDGES] = {a01.x, a01.y, a23.x, a23.y, a45.x, a45.y, a67.x, a67.y};
#pragma unroll
            for (int e = 0; e < EC_EDGES; ++e) {
                const int je = __shfl_sync(0xffffffffu, jl, e);
                if (e < nb) {
                    if (AGGR == GM_AGGR_MAX) {
                        if (acc[e] > best) { best = acc[e]; barg = je; }
                    } else {
                        sum += acc[e];
                    }
                }
            }
        }
        float o;
        const int deg = end - beg;
        if (deg == 0) o = 0.f;                                        // PyG: empty rows aggregate to 0
        else if (AGGR == GM_AGGR_MAX) o = p + best;
        else if (AGGR == GM_AGGR_ADD) o = fmaf((float)deg, p, sum);
        else o = p + sum / (float)deg;
        out[(size_t)i * H + lane] = o;
        if (ARG) arg[(size_t)i * H + lane] = barg;
        run.n += 1.f;
        const float dl = o - run.mean;
        run.mean += dl / run.n;
        run.m2 = fmaf(dl, o - run.mean, run.m2);
    }
    if (partials) {
        s_red[warp][0][lane] = run.n; s_red[warp][1][lane] = run.mean; s_red[warp][2][lane] = run.m2;
        __syncthreads();
        if (warp == 0) {
            Moments tot{0.f, 0.f, 0.f};
#pragma unroll
            for (int k = 0; k < EC_WARPS; ++k)
                tot = moments_combine(tot, Moments{s_red[k][0][lane], s_red[k][1][lane], s_red[k][2][lane]});
            float *dst = partials + (size_t)blockIdx.x * MOM_STRIDE;
            dst[lane] = tot.n; dst[H + lane] = tot.mean; dst[2 * H + lane] = tot.m2;
        }
    }
}

// ------------------------------------------------------------------------------------------ head
// emb_L = (resid +) BN(raw) folded into the load (K4), then Linear_32->16 + ELU + Linear_16->1 (+ sigmoid).
// emb_out (optional) keeps emb_L for the backward; out is written at perm[n] (NULL = identity).
__global__ void __launch_bounds__(NODE_TILE) k_head(const float *__restrict__ raw, const float *__restrict__ resid,
                                                   const float *__restrict__ stat, int N,
                                                   const float *__restrict__ w1, const float *__restrict__ b1,
                                                   const float *__restrict__ w2, const float *__restrict__ b2,
                                                   int apply_sigmoid, const int *__restrict__ perm,
                                                   float *__restrict__ emb_out, float *__restrict__ out) {
    __shared__ __align__(16) float s_w1[16 * H];
    __shared__ float s_b1[16];
    __shared__ __align__(16) float s_w2[16];
    __shared__ float s_b2[1];
    __shared__ float s_stat[BNSTAT_STRIDE];
    __shared__ float s_tile[NODE_TILE * TILE_LD];
    const int t = threadIdx.x;
    for (int i = t; i < 16 * H; i += NODE_TILE) s_w1[i] = w1[i];
    if (t < 16) { s_b1[t] = b1[t]; s_w2[t] = w2[t]; }
    if (t == 0) s_b2[0] = b2[0];
    pdl_sync();                                  // weights above: parameters; stat is the previous kernel's output
    if (stat && t < BNSTAT_STRIDE) s_stat[t] = stat[t];
    __syncthreads();
    const int tiles = (N + NODE_TILE - 1) / NODE_TILE;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const int node0 = tile * NODE_TILE, rows = min(NODE_TILE, N - node0);
        for (int idx = t; idx < rows * (H / 4); idx += NODE_TILE) {
            const int r = idx >> 3, c4 = (idx & 7) * 4;
            const size_t g = (size_t)(node0 + r) * (H / 4) + (idx & 7);
            float4 v = reinterpret_cast<const float4 *>(raw)[g];
            if (stat) {
                v.x = fmaf(v.x - s_stat[c4 + 0], s_stat[H + c4 + 0], s_stat[3 * H + c4 + 0]);
                v.y = fmaf(v.y - s_stat[c4 + 1], s_stat[H + c4 + 1], s_stat[3 * H + c4 + 1]);
                v.z = fmaf(v.z - s_stat[c4 + 2], s_stat[H + c4 + 2], s_stat[3 * H + c4 + 2]);
                v.w = fmaf(v.w - s_stat[c4 + 3], s_stat[H + c4 + 3], s_stat[3 * H + c4 + 3]);
            }
            if (resid) {
                const float4 q = reinterpret_cast<const float4 *>(resid)[g];
                v.x += q.x; v.y += q.y; v.z += q.z; v.w += q.w;
            }
            if (emb_out) reinterpret_cast<float4 *>(emb_out)[g] = v;
            float *d = s_tile + r * TILE_LD + c4;
            d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
        }
        __syncthreads();
        if (t < rows) {
            float xr[H], hdn[16], hact[16], o[1];
#pragma unroll
            for (int k = 0; k < H; ++k) xr[k] = s_tile[t * TILE_LD + k];
            dense_smem<H, 16>(s_w1, s_b1, xr, hdn);
#pragma unroll
            for (int k = 0; k < 16; ++k) hact[k] = elu(hdn[k]);
            dense_smem<16, 1>(s_w2, s_b2, hact, o);
            const int dst = perm ? perm[node0 + t] : node0 + t;
            out[dst] = apply_sigmoid ? sigmoidf(o[0]) : o[0];
        }
        __syncthreads();
    }
}

template <int AGGR>
int launch_edgeconv_simt(const float *x, const int *rowptr, const int *col, int N, const float *W, const float *b,
                         float *out, int *arg, float *partials, int *n_partials, cudaStream_t st) {
    const int grid = persistent_grid((N + EC_WARPS - 1) / EC_WARPS, 2);
    if (arg)
        k_edgeconv_simt<AGGR, true><<<grid, EC_WARPS * 32, 0, st>>>(x, rowptr, col, N, W, b, out, arg, partials);
    else
        k_edgeconv_simt<AGGR, false><<<grid, EC_WARPS * 32, 0, st>>>(x, rowptr, col, N, W, b, out, arg, partials);
    GM_LAUNCH_CHECK();
    *n_partials = grid;
    return GM_OK;
}

template <int AGGR>
int launch_edge_aggr(const float *Y, const float *P, const int *rowptr, const int *col, int N, float *out, int *arg,
                     const AggrTail &tail, cudaStream_t st) {
    // one resident wave: the grid is (#SM x resident CTAs of this instantiation), found by an occupancy query
    static int occ[2] = {0, 0};
    if (!occ[0]) occ[0] = resident_ctas(k_edge_aggr<AGGR, false>, EA_WARPS * 32, 0, 2);
    if (!occ[1]) occ[1] = resident_ctas(k_edge_aggr<AGGR, true>, EA_WARPS * 32, 0, 2);
    const int grid = persistent_grid((N + EA_WARPS - 1) / EA_WARPS, min(occ[arg ? 1 : 0], MAX_PARTIAL_CTAS / 148));
    if (arg) k_edge_aggr<AGGR, true><<<grid, EA_WARPS * 32, 0, st>>>(Y, P, rowptr, col, N, out, arg, tail);
    else k_edge_aggr<AGGR, false><<<grid, EA_WARPS * 32, 0, st>>>(Y, P, rowptr, col, N, out, arg, tail);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

int edge_aggr(int aggr, const float *Y, const float *P, const int *rowptr, const int *col, int N, float *out, int *arg,
              const AggrTail &tail, cudaStream_t st) {
    switch (aggr) {
        case GM_AGGR_MAX: return launch_edge_aggr<GM_AGGR_MAX>(Y, P, rowptr, col, N, out, arg, tail, st);
        case GM_AGGR_ADD: return launch_edge_aggr<GM_AGGR_ADD>(Y, P, rowptr, col, N, out, arg, tail, st);
        case GM_AGGR_MEAN: return launch_edge_aggr<GM_AGGR_MEAN>(Y, P, rowptr, col, N, out, arg, tail, st);
    }
    return GM_ERR_INVALID_ARGUMENT;
}

// direct per-edge contraction: fp32 CUDA cores (GM_PREC_FP32_DIRECT) or tcgen05 (tf32 / bf16)
int edgeconv_direct(const float *x, const int *rowptr, const int *col, int N, const float *W, const float *b,
                    int aggr, int precision, float *out, int *arg, float *partials, int *n_partials,
                    cudaStream_t st) {
    if (precision != GM_PREC_FP32_DIRECT)
        return edgeconv_tc_forward(x, rowptr, col, N, W, b, aggr, precision, out, arg, partials, n_partials, st);
    switch (aggr) {
        case GM_AGGR_MAX: return launch_edgeconv_simt<GM_AGGR_MAX>(x, rowptr, col, N, W, b, out, arg, partials, n_partials, st);
        case GM_AGGR_ADD: return launch_edgeconv_simt<GM_AGGR_ADD>(x, rowptr, col, N, W, b, out, arg, partials, n_partials, st);
        case GM_AGGR_MEAN: return launch_edgeconv_simt<GM_AGGR_MEAN>(x, rowptr, col, N, W, b, out, arg, partials, n_partials, st);
    }
    return GM_ERR_INVALID_ARGUMENT;
}

int node_grid(int N) { return persistent_grid((N + NODE_TILE - 1) / NODE_TILE, 4); }

}  // namespace

int check_params(const GmParams *p) {
    if (!p) return GM_ERR_INVALID_ARGUMENT;
    if (p->continuous_dim < 1 || p->continuous_dim > 8) return GM_ERR_UNSUPPORTED;
    if (p->conv_depth < 0 || p->conv_depth > GM_MAX_CONV_DEPTH) return GM_ERR_UNSUPPORTED;
    return GM_OK;
}

}  // namespace gm

using namespace gm;

extern "C" {

size_t gm_edgeconv_ws_bytes(int N) { return 2 * align_up((size_t)(N > 0 ? N : 1) * H * sizeof(float), 256); }

int gm_edgeconv_forward(const float *x, const int *rowptr, const int *col, int N, const float *W, const float *b,
                        int aggr, int precision, void *ws, size_t ws_bytes, float *out, int *arg_out, void *stream) {
    if (N < 0 || (N > 0 && (!x || !rowptr || !W || !b || !out))) return GM_ERR_INVALID_ARGUMENT;
    if (N == 0) return GM_OK;
    cudaStream_t st = as_stream(stream);
    if (precision != GM_PREC_FP32) {
        int n_part = 0;
        TimedRegion tr(GM_TIMER_EDGECONV_FWD, st);
        return edgeconv_direct(x, rowptr, col, N, W, b, aggr, precision, out, arg_out, nullptr, &n_part, st);
    }
    if (!ws) return GM_ERR_INVALID_ARGUMENT;
    if (ws_bytes < gm_edgeconv_ws_bytes(N)) return GM_ERR_CAPACITY;
    float *Y = (float *)ws;
    float *P = (float *)((char *)ws + gm_edgeconv_ws_bytes(N) / 2);
    {
        TimedRegion tr(GM_TIMER_NODE_PREP, st);
        k_node_prep2<<<node_grid(N), NP_THREADS, 0, st>>>(x, nullptr, nullptr, W, b, N, nullptr, Y, P);
        GM_LAUNCH_CHECK();
    }
    TimedRegion tr(GM_TIMER_EDGECONV_FWD, st);
    return edge_aggr(aggr, Y, P, rowptr, col, N, out, arg_out, AggrTail{}, st);
}

size_t gm_net_ws_bytes(int N, int conv_depth, int save_for_backward) {
    return net_ws_layout(N < 0 ? 0 : N, conv_depth, save_for_backward, nullptr, nullptr);
}

}  // extern "C"

// Shared body of gm_net_forward (explicit CSR) and gm_net_forward_geo (implicit radius graph: geo_ws != NULL).
static int net_forward_impl(const GmParams *p, const float *x_cont, int x_cont_stride, const int64_t *x_cat_i64,
                            const float *x_cat_f32, int x_cat_stride, const int *rowptr, const int *col,
                            const void *geo_ws, int B, int N, const int *node_perm, int training,
                            int save_for_backward, int apply_sigmoid, int aggr, int precision, float momentum,
                            float eps, void *ws, size_t ws_bytes, float *out, int *err_flags, void *stream,
                            const int *knn_ptr = nullptr, int knn_k = 0, void *knn_ws = nullptr) {
    GM_TRY(check_params(p));
    KnnLayers kl{};
    if (knn_ptr) {        // per-layer dynamic k-NN graph in the embedding space (fp32, caller node order)
        if (knn_k < 1 || knn_k > 32 || !knn_ws || B < 1 || node_perm || precision != GM_PREC_FP32 || geo_ws) return GM_ERR_UNSUPPORTED;
        knn_layers_layout(N, p->conv_depth, knn_k, (char *)knn_ws, &kl);
    }
    if (N < 0 || !ws || !err_flags) return GM_ERR_INVALID_ARGUMENT;
    if (N > 0 && (!x_cont || (!x_cat_i64 && !x_cat_f32) || (!rowptr && !geo_ws && !knn_ptr) || !out)) return GM_ERR_INVALID_ARGUMENT;
    if (geo_ws && !node_perm) return GM_ERR_UNSUPPORTED;
    // per-edge tcgen05 on the implicit graph: tf32 / bf16; training (arg-max + BatchNorm moments) for aggr = max
    if (geo_ws && precision != GM_PREC_FP32 &&
        ((precision != GM_PREC_TF32 && precision != GM_PREC_BF16) || (save_for_backward && aggr != GM_AGGR_MAX)))
        return GM_ERR_UNSUPPORTED;
    if (ws_bytes < net_ws_layout(N, p->conv_depth, save_for_backward, nullptr, nullptr)) return GM_ERR_CAPACITY;
    if (N == 0) return GM_OK;
    cudaStream_t st = as_stream(stream);
    NetWs w;
    net_ws_layout(N, p->conv_depth, save_for_backward, (char *)ws, &w);
    const int depth = p->conv_depth;
    const int ngrid = node_grid(N);

    GM_LAUNCH_PDL(k_build_lut, (LUT_ROWS * 16 + 255) / 256, 256, 0, st, *p, w.lut, w.counters, w.bnstat, training, eps);
    GM_LAUNCH_CHECK();

    EncodeArgs ea{};
    ea.x_cont = x_cont; ea.x_cont_stride = x_cont_stride;
    ea.x_cat_i64 = (const long long *)x_cat_i64;
    ea.x_cat_f32 = x_cat_f32; ea.x_cat_stride = x_cat_stride;
    ea.perm = node_perm;
    ea.N = N; ea.CD = p->continuous_dim;
    ea.flat = (!node_perm && !x_cat_i64 && x_cat_f32 && x_cat_stride == x_cont_stride && x_cont_stride <= 16 &&
               x_cat_f32 >= x_cont && (x_cat_f32 - x_cont) + 3 <= x_cont_stride) ? 1 : 0;
    ea.cont_w = p->cont_w; ea.cont_b = p->cont_b; ea.enc_w = p->enc_w; ea.enc_b = p->enc_b; ea.lut = w.lut;
    ea.enc_out = w.enc;
    ea.partials = training ? w.partials : nullptr;
    ea.counter = w.counters + 0;
    ea.bn = p->bn_all; ea.momentum = momentum; ea.eps = eps; ea.stat = w.bnstat;
    ea.err_flags = err_flags;
    {
        TimedRegion tr(GM_TIMER_ENCODE_FWD, st);
        GM_LAUNCH_PDL(k_encode2, persistent_grid((N + NODE_TILE - 1) / NODE_TILE, 3), ENC_THREADS, 0, st, ea);
        GM_LAUNCH_CHECK();
    }

    for (int l = 0; l < depth; ++l) {
        const float *raw = l == 0 ? w.enc : w.conv[l - 1];
        const float *resid = l == 0 ? nullptr : w.emb[l - 1];
        const float *stat_in = w.bnstat + (size_t)l * BNSTAT_STRIDE;
        float *stat_out = w.bnstat + (size_t)(l + 1) * BNSTAT_STRIDE;
        int *arg = save_for_backward ? w.arg[l] : nullptr;
        if (precision == GM_PREC_FP32) {
            {
                TimedRegion tr(GM_TIMER_NODE_PREP, st);
                GM_LAUNCH_PDL(k_node_prep2, ngrid, NP_THREADS, 0, st, raw, resid, stat_in, p->conv_w[l], p->conv_b[l], N,
                              w.emb[l], w.Y, w.P, 0);
                GM_LAUNCH_CHECK();
            }
            AggrTail tail{};
            if (training) tail = AggrTail{w.partials, w.counters + 1 + l, p->conv_bn[l], momentum, eps, stat_out};
            if (knn_ptr) {   // knn_graph(emb_l, k, batch, loop=True): the layer's own graph, kept for the backward
                TimedRegion trg(GM_TIMER_RADIUS_GRAPH, st);
                GM_TRY(gm_knn_graph_count(w.emb[l], H, H, knn_ptr, N, B, knn_k, 1, kl.knn_ws, kl.knn_bytes, kl.rowptr[l], stream));
                GM_TRY(gm_knn_graph_fill(N, knn_k, 1, kl.knn_ws, kl.knn_bytes, kl.rowptr[l], kl.col[l], kl.ecap, nullptr, stream));
                if (save_for_backward)
                    GM_TRY(gm_csr_transpose(kl.rowptr[l], kl.col[l], N, kl.ecap, kl.csr_ws, kl.csr_bytes, kl.rowptr_t[l],
                                            kl.col_t[l], nullptr, stream));
            }
            TimedRegion tr(GM_TIMER_EDGECONV_FWD, st);
            if (geo_ws) GM_TRY(geo_edge_aggr(aggr, w.Y, w.P, geo_ws, N, B, w.conv[l], arg, tail, st));
            else if (knn_ptr) GM_TRY(edge_aggr(aggr, w.Y, w.P, kl.rowptr[l], kl.col[l], N, w.conv[l], arg, tail, st));
            else GM_TRY(edge_aggr(aggr, w.Y, w.P, rowptr, col, N, w.conv[l], arg, tail, st));
        } else if (geo_ws) {
            {
                TimedRegion tr(GM_TIMER_NODE_PREP, st);
                GM_LAUNCH_PDL(k_node_prep2, ngrid, NP_THREADS, 0, st, raw, resid, stat_in, p->conv_w[l], p->conv_b[l], N,
                              w.emb[l], (float *)nullptr, w.P, 1);
                GM_LAUNCH_CHECK();
            }
            TimedRegion tr(GM_TIMER_EDGECONV_FWD, st);
            GM_TRY(geo_edgeconv_tc(aggr, precision, w.emb[l], w.P, p->conv_w[l], geo_ws, N, B, w.conv[l], arg, st));
            if (training)      // the tensor-core kernel is ticket-scheduled: moments as a static, bit-reproducible pass
                GM_TRY(geo_row_moments(w.conv[l], N, AggrTail{w.partials, w.counters + 1 + l, p->conv_bn[l], momentum, eps, stat_out}, st));
        } else {
            {
                TimedRegion tr(GM_TIMER_NODE_PREP, st);
                k_node_prep<false><<<ngrid, NODE_TILE, 0, st>>>(raw, resid, stat_in, nullptr, nullptr, N, w.emb[l],
                                                                nullptr, nullptr);
                GM_LAUNCH_CHECK();
            }
            int n_part = 0;
            {
                TimedRegion tr(GM_TIMER_EDGECONV_FWD, st);
                GM_TRY(edgeconv_direct(w.emb[l], rowptr, col, N, p->conv_w[l], p->conv_b[l], aggr, precision,
                                       w.conv[l], arg, training ? w.partials : nullptr, &n_part, st));
            }
            if (training) {
                k_bn_finalize<<<1, 256, 0, st>>>(w.partials, n_part, p->conv_bn[l], momentum, eps, stat_out);
                GM_LAUNCH_CHECK();
            }
        }
    }
    {
        const float *raw = depth == 0 ? w.enc : w.conv[depth - 1];
        const float *resid = depth == 0 ? nullptr : w.emb[depth - 1];
        TimedRegion tr(GM_TIMER_HEAD_FWD, st);
        GM_LAUNCH_PDL(k_head, ngrid, NODE_TILE, 0, st, raw, resid, w.bnstat + (size_t)depth * BNSTAT_STRIDE, N, p->out1_w,
                      p->out1_b, p->out2_w, p->out2_b, apply_sigmoid, node_perm,
                      save_for_backward ? w.emb[depth] : (float *)nullptr, out);
        GM_LAUNCH_CHECK();
    }
    return GM_OK;
}

extern "C" {

int gm_net_forward(const GmParams *p, const float *x_cont, int x_cont_stride, const int64_t *x_cat_i64,
                   const float *x_cat_f32, int x_cat_stride, const int *rowptr, const int *col, int N,
                   const int *node_perm, int training, int save_for_backward, int apply_sigmoid, int aggr,
                   int precision, float momentum, float eps, void *ws, size_t ws_bytes, float *out, int *err_flags,
                   void *stream) {
    return net_forward_impl(p, x_cont, x_cont_stride, x_cat_i64, x_cat_f32, x_cat_stride, rowptr, col, nullptr, 0, N,
                            node_perm, training, save_for_backward, apply_sigmoid, aggr, precision, momentum, eps, ws,
                            ws_bytes, out, err_flags, stream);
}

int gm_net_forward_geo(const GmParams *p, const float *x_cont, int x_cont_stride, const int64_t *x_cat_i64,
                       const float *x_cat_f32, int x_cat_stride, const void *geo_ws, int N, int B,
                       const int *node_perm, int training, int save_for_backward, int apply_sigmoid, int aggr,
                       int precision, float momentum, float eps, void *ws, size_t ws_bytes, float *out, int *err_flags,
                       void *stream) {
    if (!geo_ws || B < 1) return GM_ERR_INVALID_ARGUMENT;
    return net_forward_impl(p, x_cont, x_cont_stride, x_cat_i64, x_cat_f32, x_cat_stride, nullptr, nullptr, geo_ws, B, N,
                            node_perm, training, save_for_backward, apply_sigmoid, aggr, precision, momentum, eps,
                            ws, ws_bytes, out, err_flags, stream);
}

size_t gm_net_knn_ws_bytes(int N, int conv_depth, int k) {
    return knn_layers_layout(N < 0 ? 0 : N, conv_depth, k < 1 ? 1 : k, nullptr, nullptr);
}

int gm_net_forward_knn(const GmParams *p, const float *x_cont, int x_cont_stride, const int64_t *x_cat_i64,
                       const float *x_cat_f32, int x_cat_stride, const int *ptr, int N, int B, int k, int training,
                       int save_for_backward, int apply_sigmoid, int aggr, float momentum, float eps, void *ws,
                       size_t ws_bytes, void *knn_ws, size_t knn_ws_bytes, float *out, int *err_flags, void *stream) {
    if (!ptr || B < 1 || !knn_ws || !p) return GM_ERR_INVALID_ARGUMENT;
    if (knn_ws_bytes < knn_layers_layout(N, p->conv_depth, k, nullptr, nullptr)) return GM_ERR_CAPACITY;
    return net_forward_impl(p, x_cont, x_cont_stride, x_cat_i64, x_cat_f32, x_cat_stride, nullptr, nullptr, nullptr, B, N,
                            nullptr, training, save_for_backward, apply_sigmoid, aggr, GM_PREC_FP32, momentum, eps, ws,
                            ws_bytes, out, err_flags, stream, ptr, k, knn_ws);
}

int gm_edgeconv_forward_geo(const float *x, const void *geo_ws, int N, int B, const float *W, const float *b, int aggr,
                            int precision, void *ws, size_t ws_bytes, float *out, int *arg_out, void *stream) {
    if (N < 0 || B < 1 || !geo_ws || (N > 0 && (!x || !W || !b || !out || !ws))) return GM_ERR_INVALID_ARGUMENT;
    if (N == 0) return GM_OK;
    if (ws_bytes < gm_edgeconv_ws_bytes(N)) return GM_ERR_CAPACITY;
    cudaStream_t st = as_stream(stream);
    float *Y = (float *)ws;
    float *P = (float *)((char *)ws + gm_edgeconv_ws_bytes(N) / 2);
    if (precision != GM_PREC_FP32) {                    // per-edge tcgen05 contraction (arg-max for aggr = max)
        if ((arg_out && aggr != GM_AGGR_MAX) || (precision != GM_PREC_TF32 && precision != GM_PREC_BF16)) return GM_ERR_UNSUPPORTED;
        {
            TimedRegion tr(GM_TIMER_NODE_PREP, st);
            k_node_prep2<<<node_grid(N), NP_THREADS, 0, st>>>(x, nullptr, nullptr, W, b, N, nullptr, nullptr, P, 1);
            GM_LAUNCH_CHECK();
        }
        TimedRegion tr(GM_TIMER_EDGECONV_FWD, st);
        return geo_edgeconv_tc(aggr, precision, x, P, W, geo_ws, N, B, out, arg_out, st);
    }
    {
        TimedRegion tr(GM_TIMER_NODE_PREP, st);
        k_node_prep2<<<node_grid(N), NP_THREADS, 0, st>>>(x, nullptr, nullptr, W, b, N, nullptr, Y, P);
        GM_LAUNCH_CHECK();
    }
    TimedRegion tr(GM_TIMER_EDGECONV_FWD, st);
    return geo_edge_aggr(aggr, Y, P, geo_ws, N, B, out, arg_out, AggrTail{}, st);
}

}  // extern "C"
