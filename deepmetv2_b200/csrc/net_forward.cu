// Forward of the GraphMETNetwork hot path (model/graph_met_network.py:47-69, model/net.py:45-47):
//   k_build_lut   : the 168-entry categorical table ELU(Linear_24->16(cat[E_chrg,E_pdg,E_pv]))
//   k_encode      : fused embeddings + continuous encoder + encode_all (+ BatchNorm moments)   [a3,a4]
//   k_bn_finalize : deterministic combination of per-CTA moments, running-stat update        [K4]
//   k_bn_apply    : emb = (resid +) (raw - mean) * gamma*invstd + beta
//   k_edgeconv_simt: fp32 EdgeConv, warp per target, FFMA2 edge contraction, segmented max     [a5]
//   k_head        : Linear_32->16 + ELU + Linear_16->1 (+ sigmoid)                              [a6]
#include "net_layout.cuh"

namespace gm {

int edgeconv_tc_forward(const float *x, const int *rowptr, const int *col, int N, const float *W, const float *b,
                        int aggr, int precision, float *out, int *arg, float *partials, int *n_partials,
                        cudaStream_t st);

namespace {

// ------------------------------------------------------------------------------------------ LUT
__global__ void __launch_bounds__(256) k_build_lut(GmParams p, float *__restrict__ lut) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
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

// ------------------------------------------------------------------------------------------ moments
// Moments of the valid rows of a [NODE_TILE][TILE_LD] smem tile, per channel, folded into the CTA's
// running moments held by threads 0..31.  blockDim.x == NODE_TILE (4 warps).
__device__ __forceinline__ void tile_moments(const float *tile, int rows, float (*red)[3][H], Moments &run) {
    const int c = threadIdx.x & 31, q = threadIdx.x >> 5;
    const int r0 = q * 32, r1 = min(rows, r0 + 32);
    float n = (float)max(r1 - r0, 0), mean = 0.f, m2 = 0.f;
    if (r1 > r0) {
        float s = 0.f;
        for (int r = r0; r < r1; ++r) s += tile[r * TILE_LD + c];
        mean = s / n;
        for (int r = r0; r < r1; ++r) { float d = tile[r * TILE_LD + c] - mean; m2 = fmaf(d, d, m2); }
    }
    red[q][0][c] = n; red[q][1][c] = mean; red[q][2][c] = m2;
    __syncthreads();
    if (threadIdx.x < 32) {
#pragma unroll
        for (int g = 0; g < NODE_TILE / 32; ++g)
            run = moments_combine(run, Moments{red[g][0][c], red[g][1][c], red[g][2][c]});
    }
    __syncthreads();
}

__device__ __forceinline__ void store_tile(const float *tile, float *__restrict__ out, int node0, int rows) {
    // coalesced float4 stores of a [rows][32] tile
    for (int idx = threadIdx.x; idx < rows * (H / 4); idx += blockDim.x) {
        int r = idx >> 3, c4 = (idx & 7) * 4;
        const float *s = tile + r * TILE_LD + c4;
        reinterpret_cast<float4 *>(out + (size_t)(node0 + r) * H)[idx & 7] = make_float4(s[0], s[1], s[2], s[3]);
    }
}

__device__ __forceinline__ void load_tile(float *tile, const float *__restrict__ in, int node0, int rows) {
    for (int idx = threadIdx.x; idx < rows * (H / 4); idx += blockDim.x) {
        int r = idx >> 3, c4 = (idx & 7) * 4;
        float4 v = reinterpret_cast<const float4 *>(in + (size_t)(node0 + r) * H)[idx & 7];
        float *d = tile + r * TILE_LD + c4;
        d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
    }
}

// ------------------------------------------------------------------------------------------ encoder
struct EncodeArgs {
    const float *x_cont; int x_cont_stride;
    const long long *x_cat_i64;
    const float *x_cat_f32; int x_cat_stride;
    int N, CD, flat;          // flat: x_cont/x_cat_f32 are columns of one [N, stride] matrix (stride <= 16)
    const float *cont_w, *cont_b, *enc_w, *enc_b, *lut;
    float *enc_out;           // [N,32]
    float *partials;          // [grid][3][32] or NULL
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

__global__ void __launch_bounds__(NODE_TILE) k_encode(EncodeArgs a) {
    __shared__ __align__(16) float s_cont_w[16 * 8];
    __shared__ float s_cont_b[16];
    __shared__ __align__(16) float s_enc_w[H * H];
    __shared__ float s_enc_b[H];
    __shared__ __align__(16) float s_lut[LUT_ROWS * 16];
    __shared__ __align__(16) float s_x[NODE_TILE * 16];
    __shared__ float s_tile[NODE_TILE * TILE_LD];
    __shared__ float s_red[NODE_TILE / 32][3][H];

    const int t = threadIdx.x;
    for (int i = t; i < 16 * 8; i += NODE_TILE) {
        int o = i >> 3, k = i & 7;
        s_cont_w[i] = k < a.CD ? a.cont_w[o * a.CD + k] : 0.f;
    }
    if (t < 16) s_cont_b[t] = a.cont_b[t];
    for (int i = t; i < H * H; i += NODE_TILE) s_enc_w[i] = a.enc_w[i];
    if (t < H) s_enc_b[t] = a.enc_b[t];
    for (int i = t; i < LUT_ROWS * 16; i += NODE_TILE) s_lut[i] = a.lut[i];
    __syncthreads();

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
            for (int i = t; i < rows * S; i += NODE_TILE) s_x[i] = src[i];
            __syncthreads();
            if (t < rows) {
                const float *row = s_x + t * S;
                const int cat_off = (int)(a.x_cat_f32 - a.x_cont);
#pragma unroll
                for (int k = 0; k < 8; ++k) xc[k] = k < a.CD ? row[k] : 0.f;
                c0 = (long long)row[cat_off]; c1 = (long long)row[cat_off + 1]; c2 = (long long)row[cat_off + 2];
            }
        } else if (t < rows) {
            const float *row = a.x_cont + (size_t)(node0 + t) * a.x_cont_stride;
#pragma unroll
            for (int k = 0; k < 8; ++k) xc[k] = k < a.CD ? row[k] : 0.f;
            if (a.x_cat_i64) {
                const long long *cr = a.x_cat_i64 + (size_t)(node0 + t) * 3;
                c0 = cr[0]; c1 = cr[1]; c2 = cr[2];
            } else {
                const float *cr = a.x_cat_f32 + (size_t)(node0 + t) * a.x_cat_stride;
                c0 = (long long)cr[0]; c1 = (long long)cr[1]; c2 = (long long)cr[2];
            }
        }
        if (t < rows) {
            float u[H];  // cat[emb_cat, emb_cont]  (model/graph_met_network.py:58)
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
            float z[H];
            dense_smem<H, H>(s_enc_w, s_enc_b, u, z);
#pragma unroll
            for (int k = 0; k < H; ++k) s_tile[t * TILE_LD + k] = elu(z[k]);
        }
        __syncthreads();
        store_tile(s_tile, a.enc_out, node0, rows);
        if (a.partials) tile_moments(s_tile, rows, s_red, run);
        else __syncthreads();
    }
    if (a.partials && t < 32) {
        float *dst = a.partials + (size_t)blockIdx.x * MOM_STRIDE;
        dst[t] = run.n; dst[H + t] = run.mean; dst[2 * H + t] = run.m2;
    }
    if (err) atomicOr(a.err_flags, GM_FLAG_CAT_OUT_OF_RANGE);
}

// ------------------------------------------------------------------------------------------ BatchNorm
// One CTA of 1024 threads: thread (c = t%32, s = t/32) folds partial CTAs s, s+32, ... then the 32
// slices are folded in order.  Fixed order => bit-deterministic.  ATen semantics:
// biased variance for normalisation, unbiased for running_var, momentum update, counter += 1.
__global__ void __launch_bounds__(1024) k_bn_finalize(const float *__restrict__ partials, int n_partials,
                                                      GmBatchNorm bn, float momentum, float eps, int training,
                                                      float *__restrict__ stat) {
    __shared__ float red[32][3][H];
    const int c = threadIdx.x & 31, s = threadIdx.x >> 5;
    if (training) {
        Moments m{0.f, 0.f, 0.f};
#pragma unroll 4
        for (int b = s; b < n_partials; b += 32) {
            const float *p = partials + (size_t)b * MOM_STRIDE;
            m = moments_combine(m, Moments{p[c], p[H + c], p[2 * H + c]});
        }
        red[s][0][c] = m.n; red[s][1][c] = m.mean; red[s][2][c] = m.m2;
        __syncthreads();
        if (s == 0) {
            Moments tot{0.f, 0.f, 0.f};
            for (int k = 0; k < 32; ++k) tot = moments_combine(tot, Moments{red[k][0][c], red[k][1][c], red[k][2][c]});
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
    } else if (s == 0) {
        const float invstd = 1.0f / sqrtf(bn.running_var[c] + eps);
        stat[c] = bn.running_mean[c];
        stat[H + c] = bn.weight[c] * invstd;
        stat[2 * H + c] = invstd;
        stat[3 * H + c] = bn.bias[c];
    }
}

__global__ void __launch_bounds__(256) k_bn_apply(const float4 *__restrict__ raw, const float4 *__restrict__ resid,
                                                  const float *__restrict__ stat, float4 *__restrict__ out,
                                                  size_t n4) {
    __shared__ float s_stat[BNSTAT_STRIDE];
    if (threadIdx.x < BNSTAT_STRIDE) s_stat[threadIdx.x] = stat[threadIdx.x];
    __syncthreads();
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const int c = (int)(i & 7) * 4;
        float4 v = raw[i];
        float4 r = resid ? resid[i] : make_float4(0.f, 0.f, 0.f, 0.f);
        float4 o;
        o.x = r.x + fmaf(v.x - s_stat[c + 0], s_stat[H + c + 0], s_stat[3 * H + c + 0]);
        o.y = r.y + fmaf(v.y - s_stat[c + 1], s_stat[H + c + 1], s_stat[3 * H + c + 1]);
        o.z = r.z + fmaf(v.z - s_stat[c + 2], s_stat[H + c + 2], s_stat[3 * H + c + 2]);
        o.w = r.w + fmaf(v.w - s_stat[c + 3], s_stat[H + c + 3], s_stat[3 * H + c + 3]);
        out[i] = o;
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
__global__ void __launch_bounds__(NODE_TILE) k_head(const float *__restrict__ emb, int N, const float *__restrict__ w1,
                                                   const float *__restrict__ b1, const float *__restrict__ w2,
                                                   const float *__restrict__ b2, int apply_sigmoid,
                                                   float *__restrict__ out) {
    __shared__ __align__(16) float s_w1[16 * H];
    __shared__ float s_b1[16];
    __shared__ __align__(16) float s_w2[16];
    __shared__ float s_b2[1];
    __shared__ float s_tile[NODE_TILE * TILE_LD];
    const int t = threadIdx.x;
    for (int i = t; i < 16 * H; i += NODE_TILE) s_w1[i] = w1[i];
    if (t < 16) { s_b1[t] = b1[t]; s_w2[t] = w2[t]; }
    if (t == 0) s_b2[0] = b2[0];
    __syncthreads();
    const int tiles = (N + NODE_TILE - 1) / NODE_TILE;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const int node0 = tile * NODE_TILE, rows = min(NODE_TILE, N - node0);
        load_tile(s_tile, emb, node0, rows);
        __syncthreads();
        if (t < rows) {
            float xr[H], hdn[16], hact[16], o[1];
#pragma unroll
            for (int k = 0; k < H; ++k) xr[k] = s_tile[t * TILE_LD + k];
            dense_smem<H, 16>(s_w1, s_b1, xr, hdn);
#pragma unroll
            for (int k = 0; k < 16; ++k) hact[k] = elu(hdn[k]);
            dense_smem<16, 1>(s_w2, s_b2, hact, o);
            out[node0 + t] = apply_sigmoid ? sigmoidf(o[0]) : o[0];
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

}  // namespace

int edgeconv_forward(const float *x, const int *rowptr, const int *col, int N, const float *W, const float *b,
                     int aggr, int precision, float *out, int *arg, float *partials, int *n_partials,
                     cudaStream_t st) {
    if (precision != GM_PREC_FP32)
        return edgeconv_tc_forward(x, rowptr, col, N, W, b, aggr, precision, out, arg, partials, n_partials, st);
    switch (aggr) {
        case GM_AGGR_MAX: return launch_edgeconv_simt<GM_AGGR_MAX>(x, rowptr, col, N, W, b, out, arg, partials, n_partials, st);
        case GM_AGGR_ADD: return launch_edgeconv_simt<GM_AGGR_ADD>(x, rowptr, col, N, W, b, out, arg, partials, n_partials, st);
        case GM_AGGR_MEAN: return launch_edgeconv_simt<GM_AGGR_MEAN>(x, rowptr, col, N, W, b, out, arg, partials, n_partials, st);
    }
    return GM_ERR_INVALID_ARGUMENT;
}

int check_params(const GmParams *p) {
    if (!p) return GM_ERR_INVALID_ARGUMENT;
    if (p->continuous_dim < 1 || p->continuous_dim > 8) return GM_ERR_UNSUPPORTED;
    if (p->conv_depth < 0 || p->conv_depth > GM_MAX_CONV_DEPTH) return GM_ERR_UNSUPPORTED;
    return GM_OK;
}

}  // namespace gm

using namespace gm;

extern "C" {

int gm_edgeconv_forward(const float *x, const int *rowptr, const int *col, int N, const float *W, const float *b,
                        int aggr, int precision, float *out, int *arg_out, void *stream) {
    if (N < 0 || (N > 0 && (!x || !rowptr || !W || !b || !out))) return GM_ERR_INVALID_ARGUMENT;
    if (N == 0) return GM_OK;
    int n_part = 0;
    TimedRegion tr(GM_TIMER_EDGECONV_FWD, as_stream(stream));
    return edgeconv_forward(x, rowptr, col, N, W, b, aggr, precision, out, arg_out, nullptr, &n_part, as_stream(stream));
}

size_t gm_net_ws_bytes(int N, int conv_depth, int save_for_backward) {
    return net_ws_layout(N < 0 ? 0 : N, conv_depth, save_for_backward, nullptr, nullptr);
}

int gm_net_forward(const GmParams *p, const float *x_cont, int x_cont_stride, const int64_t *x_cat_i64,
                   const float *x_cat_f32, int x_cat_stride, const int *rowptr, const int *col, int N, int training,
                   int save_for_backward, int apply_sigmoid, int aggr, int precision, float momentum, float eps, void *ws, size_t ws_bytes,
                   float *out, int *err_flags, void *stream) {
    GM_TRY(check_params(p));
    if (N < 0 || !ws || !err_flags) return GM_ERR_INVALID_ARGUMENT;
    if (N > 0 && (!x_cont || (!x_cat_i64 && !x_cat_f32) || !rowptr || !out)) return GM_ERR_INVALID_ARGUMENT;
    if (ws_bytes < net_ws_layout(N, p->conv_depth, save_for_backward, nullptr, nullptr)) return GM_ERR_CAPACITY;
    if (N == 0) return GM_OK;
    cudaStream_t st = as_stream(stream);
    NetWs w;
    net_ws_layout(N, p->conv_depth, save_for_backward, (char *)ws, &w);
    const int depth = p->conv_depth;
    const int tiles = (N + NODE_TILE - 1) / NODE_TILE;

    k_build_lut<<<(LUT_ROWS * 16 + 255) / 256, 256, 0, st>>>(*p, w.lut);
    GM_LAUNCH_CHECK();

    EncodeArgs ea{};
    ea.x_cont = x_cont; ea.x_cont_stride = x_cont_stride;
    ea.x_cat_i64 = (const long long *)x_cat_i64;
    ea.x_cat_f32 = x_cat_f32; ea.x_cat_stride = x_cat_stride;
    ea.N = N; ea.CD = p->continuous_dim;
    ea.flat = (!x_cat_i64 && x_cat_f32 && x_cat_stride == x_cont_stride && x_cont_stride <= 16 &&
               x_cat_f32 >= x_cont && (x_cat_f32 - x_cont) + 3 <= x_cont_stride) ? 1 : 0;
    ea.cont_w = p->cont_w; ea.cont_b = p->cont_b; ea.enc_w = p->enc_w; ea.enc_b = p->enc_b; ea.lut = w.lut;
    ea.enc_out = w.enc;
    ea.partials = training ? w.partials : nullptr;
    ea.err_flags = err_flags;
    const int enc_grid = persistent_grid(tiles, 4);
    {
        TimedRegion tr(GM_TIMER_ENCODE_FWD, st);
        k_encode<<<enc_grid, NODE_TILE, 0, st>>>(ea);
        GM_LAUNCH_CHECK();
    }

    k_bn_finalize<<<1, 1024, 0, st>>>(w.partials, enc_grid, p->bn_all, momentum, eps, training, w.bnstat);
    GM_LAUNCH_CHECK();
    const size_t n4 = (size_t)N * (H / 4);
    const int ew_grid = persistent_grid((long long)((n4 + 255) / 256), 8);
    k_bn_apply<<<ew_grid, 256, 0, st>>>((const float4 *)w.enc, nullptr, w.bnstat, (float4 *)w.emb[0], n4);
    GM_LAUNCH_CHECK();

    for (int l = 0; l < depth; ++l) {
        int n_part = 0;
        {
            TimedRegion tr(GM_TIMER_EDGECONV_FWD, st);
            GM_TRY(edgeconv_forward(w.emb[l], rowptr, col, N, p->conv_w[l], p->conv_b[l], aggr, precision, w.conv[l],
                                    save_for_backward ? w.arg[l] : nullptr, training ? w.partials : nullptr, &n_part,
                                    st));
        }
        float *stat = w.bnstat + (size_t)(l + 1) * BNSTAT_STRIDE;
        k_bn_finalize<<<1, 1024, 0, st>>>(w.partials, n_part, p->conv_bn[l], momentum, eps, training, stat);
        GM_LAUNCH_CHECK();
        k_bn_apply<<<ew_grid, 256, 0, st>>>((const float4 *)w.conv[l], (const float4 *)w.emb[l], stat,
                                            (float4 *)w.emb[l + 1], n4);
        GM_LAUNCH_CHECK();
    }
    {
        TimedRegion tr(GM_TIMER_HEAD_FWD, st);
        k_head<<<persistent_grid(tiles, 4), NODE_TILE, 0, st>>>(w.emb[depth], N, p->out1_w, p->out1_b, p->out2_w,
                                                                p->out2_b, apply_sigmoid, out);
        GM_LAUNCH_CHECK();
    }
    return GM_OK;
}

}  // extern "C"
