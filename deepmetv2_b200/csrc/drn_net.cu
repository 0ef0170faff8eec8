// DynamicReductionNetwork forward on B200 -- /root/reference/model/dynamic_reduction_network.py:81-104
// with the wrapper of /root/reference/model/net.py:12-37.  The graph operators live in drn_graph.cu / knn.cu; this file has
// the float stages:
//   k_drn_node     datanorm * x -> inputnet (11->32->64->64, ELU) -> x0, and the per-node half of the edge MLP's first layer:
//                  Linear(128->96)(cat[x_i, x_j - x_i]) = A_i + B_j with A = (W1a - W1b) x + b1, B = W1b x  (one [N,192] row)
//   k_drn_edge     per edge (j -> i): h1 = ELU(A_i + B_j) [96]; h2 = ELU(W2 h1 + b2) [64] -- the NON-separable contraction,
//                  128-edge tiles against the K-major W2 held in shared memory; per-tile BatchNorm moments over the edge
//                  rows; per-target running max AND min (BatchNorm is a per-channel affine map, so
//                  max_j BN(h_j) = BN(max_j h_j) for a non-negative scale and BN(min_j h_j) otherwise)
//   k_drn_bn       merges the tile moments in a fixed order (train) or reads the running statistics (eval)
//   k_drn_apply    x_out[i] = BN(max or min), rows that straddle tiles folded from the per-tile head / tail partials
//   k_drn_head     global max pool per event + output MLP (64->64->32->OD) + softplus / pi(2 sigmoid - 1)
// No atomics on floats, every reduction in a fixed order: the forward is bit-deterministic run to run.
#include "geo_stage.cuh"
#include "tc_ptx.cuh"

namespace gm {

int knn_search_table(const float *x, int stride, int dim, const int *ptr, int N, int B, int K, int loop, int *nbr, int *deg,
                     const int *n_dev, cudaStream_t st);
int drn_to_undirected(const int *rowptr, const int *col, const int *nbr, const int *deg, int K, int N, const int *n_dev,
                      void *ws, size_t ws_bytes, int *rowptr_out, int *col_out, int *row_out, long long capacity,
                      long long *ei64, cudaStream_t st);
int drn_normalized_cut(const float *x, int stride, int dim, const int *rowptr, const int *col, const int *row, int N,
                       long long capacity, float *w_out, cudaStream_t st);
int drn_graclus(const int *rowptr, const int *col, const float *weight, const int *ptr, int N, int B, int men,
                const int *visit, int *cluster, int *partner, cudaStream_t st);
int drn_max_pool(const float *x, int dim, const int *cluster, const int *partner, const int *ptr, int N, int B,
                 const int *n_dev, int *scan, int *tiles, float *x_out, int *ptr_out, int *newid_out, cudaStream_t st);

namespace {

constexpr int DH = GM_DRN_HIDDEN;     // 64
constexpr int DM = GM_DRN_MIDDLE;     // 96
constexpr int DI = GM_DRN_INPUT;      // 11
constexpr int AB_W = 2 * DM;          // 192: [A | B] row per node
constexpr int DT = 128;               // nodes / edges per tile
constexpr int DTHREADS = 512;

__device__ __forceinline__ int eff_n(int N, const int *n_dev) {
    if (!n_dev) return N;
    const int v = *n_dev;
    return v < 0 ? 0 : (v < N ? v : N);
}

// global row-major W [O][K] (row pitch ldw) -> shared K-major wT [k * pitch + o0 + o]
__device__ __forceinline__ void stage_kmajor(const float *__restrict__ W, int O, int K, int ldw, float *__restrict__ wT,
                                             int pitch, int o0) {
    for (int t = threadIdx.x; t < O * K; t += blockDim.x) {
        const int o = t / K, k = t - o * K;
        wT[k * pitch + o0 + o] = W[(size_t)o * ldw + k];
    }
}

struct NodeParams {
    const float *norm, *w1, *b1, *w2, *b2, *w3, *b3;    // inputnet (INPUTNET only)
    const float *cw1, *cb1;                             // edge MLP first layer [96,128], [96]
};

// shared-memory plan (floats)
constexpr int LD_IN = 13, LD_A = 33, LD_B = 68;
constexpr int NS_WAB = DH * AB_W, NS_BAB = AB_W, NS_C = DT * LD_B;
constexpr int NS_W1 = DI * 32, NS_W2 = 32 * DH, NS_W3 = DH * DH, NS_IN = DT * LD_IN, NS_A = DT * LD_A, NS_B = DT * LD_B;
constexpr size_t NODE_SMEM_BASE = (size_t)(NS_WAB + NS_BAB + NS_C) * 4;
constexpr size_t NODE_SMEM_FULL = NODE_SMEM_BASE + (size_t)(NS_W1 + 32 + NS_W2 + DH + NS_W3 + DH + 16 + NS_IN + NS_A + NS_B) * 4;

template <bool INPUTNET>
__global__ void __launch_bounds__(DTHREADS, 1) k_drn_node(const float *__restrict__ x, int x_stride, int N,
                                                          const int *__restrict__ n_dev, NodeParams p,
                                                          float *__restrict__ x0_out, float *__restrict__ ab_out) {
    extern __shared__ __align__(16) float s_node[];
    float *sWab = s_node, *sbab = sWab + NS_WAB, *sC = sbab + NS_BAB;
    float *sW1 = sC + NS_C, *sb1 = sW1 + NS_W1, *sW2 = sb1 + 32, *sb2 = sW2 + NS_W2, *sW3 = sb2 + DH, *sb3 = sW3 + NS_W3,
          *sNorm = sb3 + DH, *sIn = sNorm + 16, *sA = sIn + NS_IN, *sB = sA + NS_A;
    const int tid = threadIdx.x;
    // ---- prologue: parameters only (overlaps the predecessor under programmatic dependent launch)
    for (int t = tid; t < DM * DH; t += DTHREADS) {       // [A | B] weights: (W1a - W1b | W1b), K-major, pitch 192
        const int o = t / DH, k = t - o * DH;
        const float wa = p.cw1[(size_t)o * 2 * DH + k], wb = p.cw1[(size_t)o * 2 * DH + DH + k];
        sWab[k * AB_W + o] = wa - wb;
        sWab[k * AB_W + DM + o] = wb;
    }
    for (int t = tid; t < AB_W; t += DTHREADS) sbab[t] = t < DM ? p.cb1[t] : 0.f;
    if constexpr (INPUTNET) {
        stage_kmajor(p.w1, 32, DI, DI, sW1, 32, 0);
        stage_kmajor(p.w2, DH, 32, 32, sW2, DH, 0);
        stage_kmajor(p.w3, DH, DH, DH, sW3, DH, 0);
        if (tid < 32) sb1[tid] = p.b1[tid];
        if (tid < DH) { sb2[tid] = p.b2[tid]; sb3[tid] = p.b3[tid]; }
        if (tid < DI) sNorm[tid] = p.norm[tid];
    }
    pdl_sync();
    const int n = eff_n(N, n_dev);
    const int ng = tid >> 4, og = tid & 15;               // tile product mapping for 64 outputs
    for (int n0 = blockIdx.x * DT; n0 < n; n0 += gridDim.x * DT) {
        const int rows = min(DT, n - n0);
        __syncthreads();                                   // previous tile fully consumed (and the prologue visible)
        if constexpr (INPUTNET) {
            for (int t = tid; t < DT * DI; t += DTHREADS) {
                const int r = t / DI, k = t - r * DI;
                sIn[r * LD_IN + k] = r < rows ? x[(size_t)(n0 + r) * x_stride + k] * sNorm[k] : 0.f;
            }
            __syncthreads();
            {   // 11 -> 32: thread = (node, 8 outputs)
                const int node = tid & (DT - 1), o0 = (tid >> 7) * 8;
                float acc[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] = sb1[o0 + j];
#pragma unroll
                for (int k = 0; k < DI; ++k) {
                    const float v = sIn[node * LD_IN + k];
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[j] = fmaf(v, sW1[k * 32 + o0 + j], acc[j]);
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) sA[node * LD_A + o0 + j] = elu(acc[j]);
            }
            __syncthreads();
            {   // 32 -> 64
                float acc[4][4];
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = sb2[4 * og + j];
                tile_gemm<32, DH, LD_A>(sA, sW2, ng, og, acc);
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    *reinterpret_cast<float4 *>(sB + (ng + 32 * i) * LD_B + 4 * og) =
                        make_float4(elu(acc[i][0]), elu(acc[i][1]), elu(acc[i][2]), elu(acc[i][3]));
            }
            __syncthreads();
            {   // 64 -> 64: x0
                float acc[4][4];
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = sb3[4 * og + j];
                tile_gemm<DH, DH, LD_B>(sB, sW3, ng, og, acc);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int r = ng + 32 * i;
                    const float4 v = make_float4(elu(acc[i][0]), elu(acc[i][1]), elu(acc[i][2]), elu(acc[i][3]));
                    *reinterpret_cast<float4 *>(sC + r * LD_B + 4 * og) = v;
                    if (r < rows) *reinterpret_cast<float4 *>(x0_out + (size_t)(n0 + r) * DH + 4 * og) = v;
                }
            }
        } else {
            for (int t = tid; t < DT * (DH / 4); t += DTHREADS) {
                const int r = t >> 4, q = t & 15;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (r < rows) v = reinterpret_cast<const float4 *>(x + (size_t)(n0 + r) * x_stride)[q];
                *reinterpret_cast<float4 *>(sC + r * LD_B + 4 * q) = v;
            }
        }
        __syncthreads();
        // [A | B] = x0 . (W1a - W1b | W1b)^T + (b1 | 0): three 64-column blocks of the 192-wide row
#pragma unroll 1
        for (int cb = 0; cb < 3; ++cb) {
            float acc[4][4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = sbab[DH * cb + 4 * og + j];
            tile_gemm<DH, DH, LD_B, AB_W>(sC, sWab + DH * cb, ng, og, acc);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int r = ng + 32 * i;
                if (r < rows)
                    *reinterpret_cast<float4 *>(ab_out + (size_t)(n0 + r) * AB_W + DH * cb + 4 * og) =
                        make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------ per-edge MLP
struct EdgeOut {
    float *mx, *mn;     // [N,64]  per-target max / min of h2 for rows that lie inside one tile
    float *bmx, *bmn;   // [2][tiles_cap][64]: 0 = head (the row started in an earlier tile), 1 = tail (continues in the next)
    float *tmom;        // [grid][3][64]: per-CTA (n, mean, M2) of h2 over the edge rows of the CTA's tiles
    int *amx, *amn;     // [N,64] source node of the max / min (first in row order on ties), NULL unless saved for the backward
    int *bamx, *bamn;   // [2][tiles_cap][64] the same for the head / tail partials
};
constexpr int LD_H1 = 100, LD_H2 = 68;
constexpr int ES_W2 = DM * DH, ES_H = DT * LD_H1, ES_MOM = 8 * DH * 2;
constexpr size_t EDGE_SMEM = (size_t)(ES_W2 + DH + ES_H + ES_MOM) * 4 + (size_t)(3 * DT + DT + 2 + 8) * 4;

// Shared tail of the per-edge kernels: the finished h2 tile ([128][LD_H2] fp32 in shared memory, rows = edges in CSR
// order) -> per-tile BatchNorm moments folded into the CTA's running moments, and per-target max / min written to the
// row's slot (inside the tile) or the tile's head / tail partial (rows that straddle tiles).
template <bool STATS, int NT>
__device__ __forceinline__ void edge_tile_reduce(const float *sH, const int *sTgt, const int *sSrc, int *sRank, int *sSeg, int *sCnt, float *sMom,
                                                 int ne, long long e0, int tile, const int *__restrict__ rowptr,
                                                 const EdgeOut &o, int tiles_cap, Moments &cta_m) {
    constexpr int PARTS = NT / DH;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid < DT) {                                       // segment starts (rows of one target are consecutive)
        const bool start = tid < ne && (tid == 0 || sTgt[tid] != sTgt[tid - 1]);
        const unsigned m = __ballot_sync(0xffffffffu, start);
        if (lane == 0) sCnt[warp] = __popc(m);
        sRank[tid] = start ? (int)__popc(m & ((1u << lane) - 1u)) : -1;        // rank of the segment inside its warp
    }
    __syncthreads();
    if (tid < DT && sRank[tid] >= 0) {
        int prefix = 0;
        for (int w = 0; w < warp; ++w) prefix += sCnt[w];
        sSeg[prefix + sRank[tid]] = tid;
    }
    const int nseg = sCnt[0] + sCnt[1] + sCnt[2] + sCnt[3];
    if (tid == 0) sSeg[nseg] = ne;
    if constexpr (STATS) {                                // per-tile BatchNorm moments, two-pass per row slice
        const int c = tid & (DH - 1), part = tid >> 6;
        float s = 0.f;
        int cnt = 0;
        for (int r = part; r < ne; r += PARTS) { s += sH[r * LD_H2 + c]; ++cnt; }
        const float mean = cnt ? s / (float)cnt : 0.f;
        float m2 = 0.f;
        for (int r = part; r < ne; r += PARTS) { const float d = sH[r * LD_H2 + c] - mean; m2 = fmaf(d, d, m2); }
        sMom[(part * DH + c) * 2] = mean;
        sMom[(part * DH + c) * 2 + 1] = m2;
    }
    __syncthreads();
    if constexpr (STATS) {
        if (tid < DH) {
            Moments m{0.f, 0.f, 0.f};
            for (int part = 0; part < PARTS; ++part) {
                const int cnt = part < ne ? (ne - part + PARTS - 1) / PARTS : 0;
                m = moments_combine(m, Moments{(float)cnt, sMom[(part * DH + tid) * 2], sMom[(part * DH + tid) * 2 + 1]});
            }
            cta_m = moments_combine(cta_m, m);
        }
    }
    for (int p = tid; p < nseg * DH; p += NT) {
        const int s = p >> 6, c = p & (DH - 1);
        const int r0 = sSeg[s], r1 = sSeg[s + 1];
        float vmax = -INFINITY, vmin = INFINITY;
        int rmax = r0, rmin = r0;
        for (int r = r0; r < r1; ++r) {                                          // strict: the first extremum of the row wins
            const float v = sH[r * LD_H2 + c];
            if (v > vmax) { vmax = v; rmax = r; }
            if (v < vmin) { vmin = v; rmin = r; }
        }
        const int i = sTgt[r0];
        const long long a = rowptr[i], b = rowptr[i + 1];
        const size_t slot = a < e0 ? ((size_t)tile) * DH + c : b > e0 + DT ? ((size_t)tiles_cap + tile) * DH + c : (size_t)i * DH + c;
        const bool partial = a < e0 || b > e0 + DT;
        (partial ? o.bmx : o.mx)[slot] = vmax;
        (partial ? o.bmn : o.mn)[slot] = vmin;
        if (o.amx) {
            (partial ? o.bamx : o.amx)[slot] = sSrc[rmax];
            (partial ? o.bamn : o.amn)[slot] = sSrc[rmin];
        }
    }
}

template <bool STATS>
__global__ void __launch_bounds__(DTHREADS, 2) k_drn_edge(const float *__restrict__ ab, const int *__restrict__ rowptr,
                                                          const int *__restrict__ row, const int *__restrict__ col, int N,
                                                          long long capacity, const float *__restrict__ w2,
                                                          const float *__restrict__ b2, EdgeOut o, int tiles_cap) {
    extern __shared__ __align__(16) float s_edge[];
    float *sW2 = s_edge, *sb2 = sW2 + ES_W2, *sH = sb2 + DH, *sMom = sH + ES_H;
    int *sTgt = reinterpret_cast<int *>(sMom + ES_MOM), *sSrc = sTgt + DT, *sRank = sSrc + DT, *sSeg = sRank + DT, *sCnt = sSeg + DT + 2;
    const int tid = threadIdx.x;
    stage_kmajor(w2, DH, DM, DM, sW2, DH, 0);
    if (tid < DH) sb2[tid] = b2[tid];
    pdl_sync();
    long long E = rowptr[N];
    if (E > capacity) E = capacity;
    const int ntiles = (int)((E + DT - 1) / DT);
    const int ng = tid >> 4, og = tid & 15;
    Moments cta_m{0.f, 0.f, 0.f};                         // threads < 64: channel tid over this CTA's tiles, fixed order
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long long e0 = (long long)tile * DT;
        const int ne = (int)min((long long)DT, E - e0);
        __syncthreads();
        if (tid < DT) {
            sTgt[tid] = tid < ne ? row[e0 + tid] : -1;
            sSrc[tid] = tid < ne ? col[e0 + tid] : -1;
        }
        __syncthreads();
        // h1 = ELU(A_i + B_j): 24 float4 per edge row
        for (int it = tid; it < DT * (DM / 4); it += DTHREADS) {
            const int e = it / (DM / 4), q = it - e * (DM / 4);
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (e < ne) {
                const float4 a = reinterpret_cast<const float4 *>(ab + (size_t)sTgt[e] * AB_W)[q];
                const float4 b = reinterpret_cast<const float4 *>(ab + (size_t)sSrc[e] * AB_W + DM)[q];
                v = make_float4(elu(a.x + b.x), elu(a.y + b.y), elu(a.z + b.z), elu(a.w + b.w));
            }
            *reinterpret_cast<float4 *>(sH + e * LD_H1 + 4 * q) = v;
        }
        __syncthreads();
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = sb2[4 * og + j];
        tile_gemm<DM, DH, LD_H1>(sH, sW2, ng, og, acc);
        __syncthreads();                                  // h1 fully consumed: the h2 tile reuses its storage
#pragma unroll
        for (int i = 0; i < 4; ++i)
            *reinterpret_cast<float4 *>(sH + (ng + 32 * i) * LD_H2 + 4 * og) =
                make_float4(elu(acc[i][0]), elu(acc[i][1]), elu(acc[i][2]), elu(acc[i][3]));
        edge_tile_reduce<STATS, DTHREADS>(sH, sTgt, sSrc, sRank, sSeg, sCnt, sMom, ne, e0, tile, rowptr, o, tiles_cap, cta_m);
    }
    if constexpr (STATS) {
        if (tid < DH) {
            o.tmom[((size_t)blockIdx.x * 3) * DH + tid] = cta_m.n;
            o.tmom[((size_t)blockIdx.x * 3 + 1) * DH + tid] = cta_m.mean;
            o.tmom[((size_t)blockIdx.x * 3 + 2) * DH + tid] = cta_m.m2;
        }
    }
}

// ------------------------------------------------------------------------------------------ per-edge MLP on tcgen05
// The same tile as k_drn_edge with the 96 -> 64 contraction on the 5th-generation tensor cores (bf16 operands, fp32
// accumulation in TMEM): this is the NON-separable edge MLP, the case a per-edge tensor-core contraction exists for.
//   A = h1 tile   [128 edges x 96] bf16, built by all threads (gather A_i + B_j, ELU, round) straight into the canonical
//       K-major no-swizzle layout (8-row x 16-byte core matrices; LBO 128 B between K chunks, SBO 1536 B between row groups)
//   B = W2        [64 x 96] bf16, staged once per CTA in the same layout
//   D = A . B^T   [128 lanes x 64 columns] fp32 in TMEM: 6 tcgen05.mma (M 128, N 64, K 16) issued by one thread, one commit
//   epilogue      8 warps: tcgen05.ld 32x32b.x32 (lane = edge row, 32 channel columns), + b2, ELU -> the fp32 h2 tile in
//                 shared memory, then the shared segmented max / min + BatchNorm-moment tail.
constexpr int TC_THREADS = 256;
constexpr int TC_LBO = 128, TC_SBO = (DM / 8) * TC_LBO;               // 12 K chunks of 8 bf16
constexpr int TC_B_BYTES = (DH / 8) * TC_SBO;      // (the A operand: (DT / 8) * TC_SBO = 24 KB)
// the fp32 h2 tile reuses the A operand's storage (the tensor core is done with A when the epilogue starts)
constexpr size_t TC_OFF_B = 0, TC_OFF_A = TC_B_BYTES, TC_OFF_H2 = TC_OFF_A, TC_OFF_MOM = TC_OFF_H2 + (size_t)DT * LD_H2 * 4,
                 TC_OFF_B2 = TC_OFF_MOM + (size_t)ES_MOM * 4, TC_OFF_INT = TC_OFF_B2 + DH * 4,
                 TC_OFF_BAR = (TC_OFF_INT + (size_t)(4 * DT + 2 + 8) * 4 + 7) / 8 * 8, TC_SMEM = TC_OFF_BAR + 16;

__device__ __forceinline__ uint64_t drn_smem_desc(uint32_t addr) {
    uint64_t d = (uint64_t)((addr >> 4) & 0x3FFF);
    d |= (uint64_t)(TC_LBO >> 4) << 16;
    d |= (uint64_t)(TC_SBO >> 4) << 32;
    d |= 1ull << 46;                                                  // descriptor version (sm_100); SWIZZLE_NONE
    return d;
}

template <bool STATS>
__global__ void __launch_bounds__(TC_THREADS, 4) k_drn_edge_tc(const float *__restrict__ ab, const int *__restrict__ rowptr,
                                                               const int *__restrict__ row, const int *__restrict__ col, int N,
                                                               long long capacity, const float *__restrict__ w2,
                                                               const float *__restrict__ b2, EdgeOut o, int tiles_cap) {
    extern __shared__ __align__(128) unsigned char s_tc[];
    float *sH2 = reinterpret_cast<float *>(s_tc + TC_OFF_H2), *sMom = reinterpret_cast<float *>(s_tc + TC_OFF_MOM),
          *sb2 = reinterpret_cast<float *>(s_tc + TC_OFF_B2);
    int *sTgt = reinterpret_cast<int *>(s_tc + TC_OFF_INT), *sSrc = sTgt + DT, *sRank = sSrc + DT, *sSeg = sRank + DT, *sCnt = sSeg + DT + 2;
    const uint32_t a_base = smem_u32(s_tc + TC_OFF_A), b_base = smem_u32(s_tc + TC_OFF_B), bar = smem_u32(s_tc + TC_OFF_BAR);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // ---- one-time setup: barrier, TMEM (64 fp32 columns), W2 as the bf16 B operand, b2
    if (tid == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) tc5_alloc(smem_u32(s_tc + TC_OFF_BAR + 8), 64);
    for (int t = tid; t < DH * (DM / 8); t += TC_THREADS) {
        const int n = t / (DM / 8), q = t - n * (DM / 8);
        const float *wp = w2 + (size_t)n * DM + 8 * q;   // scalar loads: parameters may sit at any 4-byte offset of a flat buffer
        const float4 w0 = make_float4(wp[0], wp[1], wp[2], wp[3]);
        const float4 w1 = make_float4(wp[4], wp[5], wp[6], wp[7]);
        const uint32_t addr = b_base + (uint32_t)((n >> 3) * TC_SBO + q * TC_LBO + (n & 7) * 16);
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(tc5_pack_bf16(w0.x, w0.y)),
                     "r"(tc5_pack_bf16(w0.z, w0.w)), "r"(tc5_pack_bf16(w1.x, w1.y)), "r"(tc5_pack_bf16(w1.z, w1.w)) : "memory");
    }
    if (tid < DH) sb2[tid] = b2[tid];
    tc5_fence_async_smem();
    tc5_fence_before();
    __syncthreads();
    tc5_fence_after();
    pdl_sync();
    const uint32_t tmem = *reinterpret_cast<const uint32_t *>(s_tc + TC_OFF_BAR + 8);
    // instruction descriptor: fp32 accumulate, bf16 A and B, both K-major, N = 64, M = 128
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(DH >> 3) << 17) | ((uint32_t)(DT >> 4) << 24);
    const uint64_t adesc0 = drn_smem_desc(a_base), bdesc0 = drn_smem_desc(b_base);
    long long E = rowptr[N];
    if (E > capacity) E = capacity;
    const int ntiles = (int)((E + DT - 1) / DT);
    Moments cta_m{0.f, 0.f, 0.f};
    uint32_t parity = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long long e0 = (long long)tile * DT;
        const int ne = (int)min((long long)DT, E - e0);
        if (tid < DT) {
            sTgt[tid] = tid < ne ? row[e0 + tid] : -1;
            sSrc[tid] = tid < ne ? col[e0 + tid] : -1;
        }
        __syncthreads();
        // A operand: h1 = ELU(A_i + B_j) rounded to bf16, 8 values (one 16-byte K chunk of one row) per item
        for (int it = tid; it < DT * (DM / 8); it += TC_THREADS) {
            const int e = it / (DM / 8), q = it - e * (DM / 8);
            uint32_t p0 = 0, p1 = 0, p2 = 0, p3 = 0;
            if (e < ne) {
                const float4 *pa = reinterpret_cast<const float4 *>(ab + (size_t)sTgt[e] * AB_W + 8 * q);
                const float4 *pb = reinterpret_cast<const float4 *>(ab + (size_t)sSrc[e] * AB_W + DM + 8 * q);
                const float4 a0 = pa[0], a1 = pa[1], c0 = pb[0], c1 = pb[1];
                p0 = tc5_pack_bf16(elu_fast(a0.x + c0.x), elu_fast(a0.y + c0.y));
                p1 = tc5_pack_bf16(elu_fast(a0.z + c0.z), elu_fast(a0.w + c0.w));
                p2 = tc5_pack_bf16(elu_fast(a1.x + c1.x), elu_fast(a1.y + c1.y));
                p3 = tc5_pack_bf16(elu_fast(a1.z + c1.z), elu_fast(a1.w + c1.w));
            }
            const uint32_t addr = a_base + (uint32_t)((e >> 3) * TC_SBO + q * TC_LBO + (e & 7) * 16);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(p0), "r"(p1), "r"(p2), "r"(p3) : "memory");
        }
        tc5_fence_async_smem();                           // generic-proxy stores -> visible to the tensor core's async proxy
        tc5_fence_before();
        __syncthreads();
        if (tid == 0) {
            tc5_fence_after();
#pragma unroll
            for (int ks = 0; ks < DM / 16; ++ks)
                tc5_mma<false>(tmem, adesc0 + (uint64_t)((ks * 2 * TC_LBO) >> 4), bdesc0 + (uint64_t)((ks * 2 * TC_LBO) >> 4), idesc,
                               ks > 0 ? 1u : 0u);
            tc5_commit(bar);
        }
        mbar_wait(bar, parity, 32);
        parity ^= 1u;
        tc5_fence_after();
        {   // epilogue: warp w reads lane quarter (w & 3), channel columns 32 * (w >> 2) .. + 31
            const int quarter = warp & 3, cb = warp >> 2;
            uint32_t r[32];
            tc5_ld32(tmem + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(cb * 32), r);
            float *dst = sH2 + (quarter * 32 + lane) * LD_H2 + cb * 32;
#pragma unroll
            for (int c = 0; c < 32; c += 4)
                *reinterpret_cast<float4 *>(dst + c) =
                    make_float4(elu_fast(__uint_as_float(r[c]) + sb2[cb * 32 + c]), elu_fast(__uint_as_float(r[c + 1]) + sb2[cb * 32 + c + 1]),
                                elu_fast(__uint_as_float(r[c + 2]) + sb2[cb * 32 + c + 2]), elu_fast(__uint_as_float(r[c + 3]) + sb2[cb * 32 + c + 3]));
        }
        tc5_fence_before();
        __syncthreads();                                  // h2 complete; TMEM and the A tile are free for the next tile
        edge_tile_reduce<STATS, TC_THREADS>(sH2, sTgt, sSrc, sRank, sSeg, sCnt, sMom, ne, e0, tile, rowptr, o, tiles_cap, cta_m);
        __syncthreads();
    }
    if constexpr (STATS) {
        if (tid < DH) {
            o.tmom[((size_t)blockIdx.x * 3) * DH + tid] = cta_m.n;
            o.tmom[((size_t)blockIdx.x * 3 + 1) * DH + tid] = cta_m.mean;
            o.tmom[((size_t)blockIdx.x * 3 + 2) * DH + tid] = cta_m.m2;
        }
    }
    tc5_fence_before();
    __syncthreads();
    tc5_fence_after();
    if (warp == 0) tc5_dealloc(tmem, 64);
}

// bnss [4][64] = mean, invstd, gamma, beta of the BatchNorm1d over the edge rows.  Training: the per-CTA Chan moments
// of k_drn_edge (ncta of them) are merged in a fixed two-level order (deterministic).
__global__ void __launch_bounds__(1024) k_drn_bn(const int *__restrict__ rowptr, int N, long long capacity,
                                                 const float *__restrict__ tmom, int ncta, int training, float momentum,
                                                 float eps, const float *__restrict__ gamma, const float *__restrict__ beta,
                                                 float *__restrict__ running_mean, float *__restrict__ running_var,
                                                 long long *__restrict__ nbt, float *__restrict__ bnss) {
    __shared__ float s_m[16][DH][3];
    pdl_sync();
    const int tid = threadIdx.x, c = tid & (DH - 1), part = tid >> 6;
    long long E = rowptr[N];
    if (E > capacity) E = capacity;
    const bool batch_stats = training && E > 0;
    if (batch_stats) {
        Moments m{0.f, 0.f, 0.f};
        for (int t = part; t < ncta; t += 16)
            m = moments_combine(m, Moments{tmom[((size_t)t * 3) * DH + c], tmom[((size_t)t * 3 + 1) * DH + c],
                                           tmom[((size_t)t * 3 + 2) * DH + c]});
        s_m[part][c][0] = m.n; s_m[part][c][1] = m.mean; s_m[part][c][2] = m.m2;
    }
    __syncthreads();
    if (tid >= DH) return;
    float mean, invstd;
    if (batch_stats) {
        Moments m{0.f, 0.f, 0.f};
        for (int q = 0; q < 16; ++q) m = moments_combine(m, Moments{s_m[q][c][0], s_m[q][c][1], s_m[q][c][2]});
        mean = m.mean;
        const float var = m.m2 / m.n;
        invstd = 1.f / sqrtf(var + eps);
        const float unbiased = m.n > 1.f ? m.m2 / (m.n - 1.f) : var;
        running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
        running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
        if (c == 0 && nbt) nbt[0] += 1;
    } else {
        mean = running_mean[c];
        invstd = 1.f / sqrtf(running_var[c] + eps);
    }
    bnss[c] = mean; bnss[DH + c] = invstd; bnss[2 * DH + c] = gamma[c]; bnss[3 * DH + c] = beta[c];
}

__global__ void __launch_bounds__(256) k_drn_apply(const int *__restrict__ rowptr, int N, const int *__restrict__ n_dev,
                                                   EdgeOut o, int tiles_cap, const float *__restrict__ bnss,
                                                   float *__restrict__ x_out, float *__restrict__ vsel_out,
                                                   int *__restrict__ asel_out) {
    pdl_sync();
    const int n = eff_n(N, n_dev);
    const long long total = (long long)n * DH;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        const int i = (int)(t >> 6), c = (int)(t & (DH - 1));
        const int a = rowptr[i], b = rowptr[i + 1];
        float out = 0.f, vs = 0.f;                        // no incoming edge: scatter-max leaves 0
        int as = -1;
        if (b > a) {
            const int t0 = a >> 7, t1 = (b - 1) >> 7;
            float vmax, vmin;
            int amax = -1, amin = -1;
            if (t0 == t1) {
                vmax = o.mx[(size_t)i * DH + c];
                vmin = o.mn[(size_t)i * DH + c];
                if (o.amx) { amax = o.amx[(size_t)i * DH + c]; amin = o.amn[(size_t)i * DH + c]; }
            } else {
                const size_t s0 = ((size_t)tiles_cap + t0) * DH + c;
                vmax = o.bmx[s0];
                vmin = o.bmn[s0];
                if (o.amx) { amax = o.bamx[s0]; amin = o.bamn[s0]; }
                for (int q = t0 + 1; q <= t1; ++q) {      // strict comparisons in row order: the first extremum wins
                    const size_t sq = (size_t)q * DH + c;
                    const float hx = o.bmx[sq], hn = o.bmn[sq];
                    if (hx > vmax) { vmax = hx; if (o.amx) amax = o.bamx[sq]; }
                    if (hn < vmin) { vmin = hn; if (o.amx) amin = o.bamn[sq]; }
                }
            }
            const float g = bnss[2 * DH + c];
            const float v = g >= 0.f ? vmax : vmin;       // BatchNorm is monotone per channel: max commutes up to the sign
            out = (v - bnss[c]) * bnss[DH + c] * g + bnss[3 * DH + c];
            vs = v;
            as = g >= 0.f ? amax : amin;
        }
        x_out[t] = out;
        if (vsel_out) vsel_out[t] = vs;                   // the selected (max or min) h2 and the source node it came from:
        if (asel_out) asel_out[t] = as;                   // what the backward routes the gradient through
    }
}

// CTA per event: global max pool over the event's (pooled) nodes, then the output MLP and the wrapper's activations
struct HeadParams {
    const float *w1, *b1, *w2, *b2, *w3, *b3;
    int od;
};
constexpr int HEAD_PARTS = 16;
__global__ void __launch_bounds__(HEAD_PARTS * DH) k_drn_head(const float *__restrict__ x, const int *__restrict__ ptr, int B, HeadParams p,
                                                  float *__restrict__ pooled, float *__restrict__ out,
                                                  float *__restrict__ net_out) {
    __shared__ float s_part[HEAD_PARTS][DH], s_p[DH], s_h1[DH], s_h2[DH / 2], s_o[8];
    pdl_sync();
    const int tid = threadIdx.x, c = tid & (DH - 1), part = tid >> 6;
    for (int e = blockIdx.x; e < B; e += gridDim.x) {
        const int a = ptr[e], b = ptr[e + 1];
        float m = -INFINITY;
#pragma unroll 4
        for (int r = a + part; r < b; r += HEAD_PARTS) m = fmaxf(m, x[(size_t)r * DH + c]);
        s_part[part][c] = m;
        __syncthreads();
        if (tid < DH) {
            float v = s_part[0][tid];
#pragma unroll
            for (int q = 1; q < HEAD_PARTS; ++q) v = fmaxf(v, s_part[q][tid]);
            if (b <= a) v = 0.f;
            s_p[tid] = v;
            if (pooled) pooled[(size_t)e * DH + tid] = v;
        }
        __syncthreads();
        if (tid < DH) {
            float acc = p.b1[tid];
            for (int k = 0; k < DH; ++k) acc = fmaf(p.w1[tid * DH + k], s_p[k], acc);
            s_h1[tid] = elu(acc);
        }
        __syncthreads();
        if (tid < DH / 2) {
            float acc = p.b2[tid];
            for (int k = 0; k < DH; ++k) acc = fmaf(p.w2[tid * DH + k], s_h1[k], acc);
            s_h2[tid] = elu(acc);
        }
        __syncthreads();
        if (tid < p.od) {
            float acc = p.b3[tid];
            for (int k = 0; k < DH / 2; ++k) acc = fmaf(p.w3[tid * (DH / 2) + k], s_h2[k], acc);
            s_o[tid] = acc;
            out[(size_t)e * p.od + tid] = acc;
        }
        __syncthreads();
        if (net_out && tid == 0 && p.od == 2) {           // model/net.py:31-36
            const float o0 = s_o[0], o1 = s_o[1];
            net_out[2 * e] = o0 > 20.f ? o0 : log1pf(expf(o0));
            net_out[2 * e + 1] = 3.14159265358979323846f * (2.f * sigmoidf(o1) - 1.f);
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------ workspace
constexpr int MAX_EDGE_CTAS = 148 * 4;
struct DrnWs {
    float *x0, *ab, *mx, *mn, *bmx, *bmn, *tmom, *bnss, *x1, *weight1, *x1p, *x2, *pooled, *vsel1, *vsel2;
    int *amx, *amn, *bamx, *bamn, *asel1, *asel2;
    int *nbr, *deg, *rowptr1, *col1, *row1, *cluster1, *partner1, *scan, *scan_tiles, *ptr1p, *rowptr2, *col2, *row2;
    char *und;
    size_t und_bytes;
    long long ecap;
    int tiles_cap;
};

size_t drn_layout(int N, int B, int k, char *base, DrnWs *w) {
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return base ? base + o : (char *)nullptr; };
    const size_t n = (size_t)(N > 0 ? N : 1), K = (size_t)k + 1;
    DrnWs t{};
    t.ecap = (long long)(2 * n * K);
    t.tiles_cap = (int)(t.ecap / DT + 2);
    t.x0 = (float *)take(n * DH * 4);
    t.ab = (float *)take(n * AB_W * 4);
    t.mx = (float *)take(n * DH * 4);
    t.mn = (float *)take(n * DH * 4);
    t.bmx = (float *)take((size_t)2 * t.tiles_cap * DH * 4);
    t.bmn = (float *)take((size_t)2 * t.tiles_cap * DH * 4);
    t.tmom = (float *)take((size_t)3 * MAX_EDGE_CTAS * DH * 4);
    t.bnss = (float *)take(2 * 4 * DH * 4);
    t.x1 = (float *)take(n * DH * 4);
    t.weight1 = (float *)take((size_t)t.ecap * 4);
    t.x1p = (float *)take(n * DH * 4);
    t.x2 = (float *)take(n * DH * 4);
    t.pooled = (float *)take((size_t)(B > 0 ? B : 1) * DH * 4);
    t.vsel1 = (float *)take(n * DH * 4);
    t.vsel2 = (float *)take(n * DH * 4);
    t.amx = (int *)take(n * DH * 4);
    t.amn = (int *)take(n * DH * 4);
    t.bamx = (int *)take((size_t)2 * t.tiles_cap * DH * 4);
    t.bamn = (int *)take((size_t)2 * t.tiles_cap * DH * 4);
    t.asel1 = (int *)take(n * DH * 4);
    t.asel2 = (int *)take(n * DH * 4);
    t.nbr = (int *)take(n * K * 4);
    t.deg = (int *)take((n + 1) * 4);
    t.rowptr1 = (int *)take((n + 1) * 4);
    t.col1 = (int *)take((size_t)t.ecap * 4);
    t.row1 = (int *)take((size_t)t.ecap * 4);
    t.cluster1 = (int *)take(n * 4);
    t.partner1 = (int *)take(n * 4);
    t.scan = (int *)take((n + 1) * 4);
    t.scan_tiles = (int *)take(((n + 1) / 1024 + 2) * 4);
    t.ptr1p = (int *)take((size_t)(B + 2) * 4);
    t.rowptr2 = (int *)take((n + 1) * 4);
    t.col2 = (int *)take((size_t)t.ecap * 4);
    t.row2 = (int *)take((size_t)t.ecap * 4);
    t.und_bytes = gm_undirected_ws_bytes((int)n, (long long)(n * K));
    t.und = take(t.und_bytes);
    if (w) *w = t;
    return off;
}

int edge_stage(const DrnWs &w, const GmDrnConv &cv, const int *rowptr, const int *col, const int *row, int N, const int *n_dev,
               int training, int precision, float momentum, float eps, float *bnss, float *x_out, float *vsel_out,
               int *asel_out, cudaStream_t st) {
    EdgeOut o{w.mx, w.mn, w.bmx, w.bmn, w.tmom, nullptr, nullptr, nullptr, nullptr};
    if (training) { o.amx = w.amx; o.amn = w.amn; o.bamx = w.bamx; o.bamn = w.bamn; }
    static bool attr = false;
    if (!attr) {
        GM_CUDA_TRY(cudaFuncSetAttribute(k_drn_edge<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)EDGE_SMEM));
        GM_CUDA_TRY(cudaFuncSetAttribute(k_drn_edge<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)EDGE_SMEM));
        GM_CUDA_TRY(cudaFuncSetAttribute(k_drn_edge_tc<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM));
        GM_CUDA_TRY(cudaFuncSetAttribute(k_drn_edge_tc<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM));
        attr = true;
    }
    const bool tc = precision == GM_PREC_BF16;
    int grid = persistent_grid(w.tiles_cap, tc ? 4 : 2);
    if (grid > MAX_EDGE_CTAS) grid = MAX_EDGE_CTAS;
    {
    TimedRegion tr(GM_TIMER_DRN_EDGE, st);
    if (tc) {
        if (training)
            GM_LAUNCH_PDL(k_drn_edge_tc<true>, grid, TC_THREADS, TC_SMEM, st, (const float *)w.ab, rowptr, row, col, N, w.ecap,
                          (const float *)cv.w2, (const float *)cv.b2, o, w.tiles_cap);
        else
            GM_LAUNCH_PDL(k_drn_edge_tc<false>, grid, TC_THREADS, TC_SMEM, st, (const float *)w.ab, rowptr, row, col, N, w.ecap,
                          (const float *)cv.w2, (const float *)cv.b2, o, w.tiles_cap);
    } else if (training) {
        GM_LAUNCH_PDL(k_drn_edge<true>, grid, DTHREADS, EDGE_SMEM, st, (const float *)w.ab, rowptr, row, col, N, w.ecap,
                      (const float *)cv.w2, (const float *)cv.b2, o, w.tiles_cap);
    } else {
        GM_LAUNCH_PDL(k_drn_edge<false>, grid, DTHREADS, EDGE_SMEM, st, (const float *)w.ab, rowptr, row, col, N, w.ecap,
                      (const float *)cv.w2, (const float *)cv.b2, o, w.tiles_cap);
    }
    GM_LAUNCH_CHECK();
    }
    TimedRegion tr2(GM_TIMER_DRN_POOL, st);
    GM_LAUNCH_PDL(k_drn_bn, 1, 1024, 0, st, rowptr, N, w.ecap, (const float *)w.tmom, grid, training, momentum, eps,
                  (const float *)cv.bn_weight, (const float *)cv.bn_bias, cv.bn_running_mean, cv.bn_running_var,
                  (long long *)cv.bn_num_batches_tracked, bnss);
    GM_LAUNCH_CHECK();
    GM_LAUNCH_PDL(k_drn_apply, persistent_grid(((long long)N * DH + 255) / 256, 8), 256, 0, st, rowptr, N, n_dev, o,
                  w.tiles_cap, (const float *)bnss, x_out, vsel_out, training ? asel_out : (int *)nullptr);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

// ====================================================================================================== backward
// Backward of gm_drn_forward (training, fp32): d out -> gradients of all 31 tensors.  Max / pool routing is recovered
// by EQUALITY with the stored selections (the kernels recompute h2 with the forward's own tile code, bit for bit),
// BatchNorm over the edge rows is differentiated with batch statistics (dense over all edges), parameter gradients
// are per-CTA partials summed in a fixed order: no float atomics, bit-deterministic.

// dst[j] = sum_r src[r * len + j], rows ascending
__global__ void __launch_bounds__(256) k_sum_rows(const float *__restrict__ src, int rows, int len, float *__restrict__ dst) {
    pdl_sync();
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < len; j += gridDim.x * blockDim.x) {
        float acc = 0.f;
        for (int r = 0; r < rows; ++r) acc += src[(size_t)r * len + j];
        dst[j] = acc;
    }
}

// head: recompute the output MLP of one event, back-propagate d out, route d pooled to the arg-max rows of x2
constexpr int HB_LEN = DH * DH + DH + (DH / 2) * DH + DH / 2 + 8 * (DH / 2) + 8;     // per-event parameter-gradient partial
__global__ void __launch_bounds__(256) k_drn_head_bwd(const float *__restrict__ x2, const int *__restrict__ ptr, int B,
                                                      const float *__restrict__ pooled, HeadParams p,
                                                      const float *__restrict__ dout, float *__restrict__ part,
                                                      float *__restrict__ dx2) {
    __shared__ float s_p[DH], s_h1[DH], s_h2[DH / 2], s_do[8], s_dz2[DH / 2], s_dz1[DH], s_dp[DH];
    __shared__ int s_first[4][DH];
    pdl_sync();
    const int tid = threadIdx.x;
    for (int e = blockIdx.x; e < B; e += gridDim.x) {
        float *out = part + (size_t)e * HB_LEN;
        if (tid < DH) s_p[tid] = pooled[(size_t)e * DH + tid];
        if (tid < p.od) s_do[tid] = dout[(size_t)e * p.od + tid];
        __syncthreads();
        if (tid < DH) {
            float acc = p.b1[tid];
            for (int k = 0; k < DH; ++k) acc = fmaf(p.w1[tid * DH + k], s_p[k], acc);
            s_h1[tid] = elu(acc);
        }
        __syncthreads();
        if (tid < DH / 2) {
            float acc = p.b2[tid];
            for (int k = 0; k < DH; ++k) acc = fmaf(p.w2[tid * DH + k], s_h1[k], acc);
            s_h2[tid] = elu(acc);
        }
        __syncthreads();
        // layer 3: dW3 = do (x) h2, db3 = do, dh2 = W3^T do
        for (int t = tid; t < 8 * (DH / 2); t += 256) {
            const int o = t / (DH / 2), k = t - o * (DH / 2);
            out[DH * DH + DH + (DH / 2) * DH + DH / 2 + t] = o < p.od ? s_do[o] * s_h2[k] : 0.f;
        }
        if (tid < 8) out[DH * DH + DH + (DH / 2) * DH + DH / 2 + 8 * (DH / 2) + tid] = tid < p.od ? s_do[tid] : 0.f;
        if (tid < DH / 2) {
            float acc = 0.f;
            for (int o = 0; o < p.od; ++o) acc = fmaf(p.w3[o * (DH / 2) + tid], s_do[o], acc);
            s_dz2[tid] = acc * elu_grad_from_out(s_h2[tid]);
        }
        __syncthreads();
        for (int t = tid; t < (DH / 2) * DH; t += 256) {
            const int o = t / DH, k = t - o * DH;
            out[DH * DH + DH + t] = s_dz2[o] * s_h1[k];
        }
        if (tid < DH / 2) out[DH * DH + DH + (DH / 2) * DH + tid] = s_dz2[tid];
        if (tid < DH) {
            float acc = 0.f;
            for (int o = 0; o < DH / 2; ++o) acc = fmaf(p.w2[o * DH + tid], s_dz2[o], acc);
            s_dz1[tid] = acc * elu_grad_from_out(s_h1[tid]);
        }
        __syncthreads();
        for (int t = tid; t < DH * DH; t += 256) {
            const int o = t / DH, k = t - o * DH;
            out[t] = s_dz1[o] * s_p[k];
        }
        if (tid < DH) {
            out[DH * DH + tid] = s_dz1[tid];
            float acc = 0.f;
            for (int o = 0; o < DH; ++o) acc = fmaf(p.w1[o * DH + tid], s_dz1[o], acc);
            s_dp[tid] = acc;
        }
        // global max pool: the gradient of channel c goes to the FIRST row of the event that attains the maximum
        const int a = ptr[e], b = ptr[e + 1], c = tid & (DH - 1), part4 = tid >> 6;
        int first = 0x7fffffff;
        for (int r = a + part4; r < b; r += 4)
            if (x2[(size_t)r * DH + c] == s_p[c]) { first = r; break; }
        s_first[part4][c] = first;
        __syncthreads();
        if (tid < DH) s_first[0][tid] = min(min(s_first[0][tid], s_first[1][tid]), min(s_first[2][tid], s_first[3][tid]));
        __syncthreads();
        for (long long t = tid; t < (long long)(b - a) * DH; t += 256) {
            const int r = a + (int)(t >> 6), cc = (int)(t & (DH - 1));
            dx2[(size_t)r * DH + cc] = r == s_first[0][cc] ? s_dp[cc] : 0.f;
        }
        __syncthreads();
    }
}

// S_c = sum_i g[i,c], T_c = sum_i g[i,c] * hat_sel(i,c) over the nodes that have edges (g = d x_out); also d gamma, d beta.
// Two launches: per-CTA partials over an interleaved node range (fixed order inside a CTA), then their sum in CTA order.
constexpr int BS_PARTS = 16;
__global__ void __launch_bounds__(BS_PARTS * DH) k_drn_bn_sums(const float *__restrict__ g, const float *__restrict__ vsel,
                                                               const int *__restrict__ rowptr, int N, const int *__restrict__ n_dev,
                                                               const float *__restrict__ bnss, double *__restrict__ partial) {
    __shared__ double s_s[BS_PARTS][DH], s_t[BS_PARTS][DH];
    pdl_sync();
    const int n = eff_n(N, n_dev), tid = threadIdx.x, c = tid & (DH - 1), part = tid >> 6;
    const float mean = bnss[c], invstd = bnss[DH + c];
    double S = 0.0, T = 0.0;
    for (int i = blockIdx.x * BS_PARTS + part; i < n; i += gridDim.x * BS_PARTS) {
        if (rowptr[i + 1] <= rowptr[i]) continue;
        const float gi = g[(size_t)i * DH + c];
        S += (double)gi;
        T += (double)(gi * ((vsel[(size_t)i * DH + c] - mean) * invstd));
    }
    s_s[part][c] = S; s_t[part][c] = T;
    __syncthreads();
    if (tid < DH) {
        double a = 0.0, b = 0.0;
        for (int q = 0; q < BS_PARTS; ++q) { a += s_s[q][tid]; b += s_t[q][tid]; }
        partial[((size_t)blockIdx.x * 2) * DH + tid] = a;
        partial[((size_t)blockIdx.x * 2 + 1) * DH + tid] = b;
    }
}
__global__ void __launch_bounds__(2 * DH) k_drn_bn_sums_fin(const double *__restrict__ partial, int nparts, float *__restrict__ sums,
                                                            float *__restrict__ dgamma, float *__restrict__ dbeta) {
    pdl_sync();
    const int tid = threadIdx.x;                       // 0..63: S, 64..127: T
    double acc = 0.0;
    for (int q = 0; q < nparts; ++q) acc += partial[(size_t)q * 2 * DH + tid];
    sums[tid] = (float)acc;
    if (tid < DH) { if (dbeta) dbeta[tid] = (float)acc; }
    else if (dgamma) dgamma[tid - DH] = (float)acc;
}

// position of the reverse entry: mirror[p] = index of (col -> row) for entry p = (row -> col) of a symmetric sorted CSR
__global__ void __launch_bounds__(256) k_drn_mirror(const int *__restrict__ rowptr, const int *__restrict__ col,
                                                    const int *__restrict__ row, int N, long long capacity, int *__restrict__ mirror) {
    pdl_sync();
    long long E = rowptr[N];
    if (E > capacity) E = capacity;
    for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < E; p += (long long)gridDim.x * blockDim.x) {
        const int r = row[p], c = col[p];
        int lo = rowptr[c], hi = rowptr[c + 1] - 1;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (col[mid] < r) lo = mid + 1; else hi = mid;
        }
        mirror[p] = lo;
    }
}

// per 128-edge tile: recompute h1, h2 (the forward's tile code), BatchNorm + max backward -> dz2; per-CTA dW2 / db2;
// dz1 = (W2^T dz2) * ELU'(h1) written per edge (the node side gathers it by target and, through `mirror`, by source)
constexpr int EB_LD_DZ = 68;
constexpr size_t EB_SMEM = (size_t)(2 * DM * DH + DH + DT * LD_H1 + DT * EB_LD_DZ + 6 * DH) * 4 + (size_t)(2 * DT) * 4;
constexpr int EB_PART = DH * DM + DH;            // dW2 [64][96] | db2 [64]
__global__ void __launch_bounds__(DTHREADS, 1) k_drn_edge_bwd(const float *__restrict__ ab, const int *__restrict__ rowptr,
                                                              const int *__restrict__ row, const int *__restrict__ col, int N,
                                                              long long capacity, const float *__restrict__ w2,
                                                              const float *__restrict__ b2, const float *__restrict__ bnss,
                                                              const float *__restrict__ sums, const int *__restrict__ asel,
                                                              const float *__restrict__ gout, float *__restrict__ dz1,
                                                              float *__restrict__ part) {
    extern __shared__ __align__(16) float s_eb[];
    float *sW2T = s_eb, *sW2 = sW2T + DM * DH, *sb2 = sW2 + DM * DH, *sH1 = sb2 + DH, *sDZ = sH1 + DT * LD_H1,
          *sBn = sDZ + DT * EB_LD_DZ;
    int *sTgt = reinterpret_cast<int *>(sBn + 6 * DH), *sSrc = sTgt + DT;
    const int tid = threadIdx.x;
    stage_kmajor(w2, DH, DM, DM, sW2T, DH, 0);
    for (int t = tid; t < DH * DM; t += DTHREADS) sW2[t] = w2[t];                   // row-major [c][k]: the W2^T dz2 product
    if (tid < DH) sb2[tid] = b2[tid];
    pdl_sync();
    long long E = rowptr[N];
    if (E > capacity) E = capacity;
    if (tid < DH) {
        sBn[tid] = bnss[tid]; sBn[DH + tid] = bnss[DH + tid]; sBn[2 * DH + tid] = bnss[2 * DH + tid];
        sBn[3 * DH + tid] = E > 0 ? sums[tid] / (float)E : 0.f;                     // S / E
        sBn[4 * DH + tid] = E > 0 ? sums[DH + tid] / (float)E : 0.f;                // T / E
    }
    const int ntiles = (int)((E + DT - 1) / DT);
    const int ng = tid >> 4, og = tid & 15;
    float accW[4][3], accb = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) accW[i][j] = 0.f;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long long e0 = (long long)tile * DT;
        const int ne = (int)min((long long)DT, E - e0);
        __syncthreads();
        if (tid < DT) {
            sTgt[tid] = tid < ne ? row[e0 + tid] : -1;
            sSrc[tid] = tid < ne ? col[e0 + tid] : -1;
        }
        __syncthreads();
        for (int it = tid; it < DT * (DM / 4); it += DTHREADS) {                    // h1: identical to k_drn_edge
            const int e = it / (DM / 4), q = it - e * (DM / 4);
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (e < ne) {
                const float4 a = reinterpret_cast<const float4 *>(ab + (size_t)sTgt[e] * AB_W)[q];
                const float4 b = reinterpret_cast<const float4 *>(ab + (size_t)sSrc[e] * AB_W + DM)[q];
                v = make_float4(elu(a.x + b.x), elu(a.y + b.y), elu(a.z + b.z), elu(a.w + b.w));
            }
            *reinterpret_cast<float4 *>(sH1 + e * LD_H1 + 4 * q) = v;
        }
        __syncthreads();
        {
            float acc[4][4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = sb2[4 * og + j];
            tile_gemm<DM, DH, LD_H1>(sH1, sW2T, ng, og, acc);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int e = ng + 32 * i;
                float4 dz = make_float4(0.f, 0.f, 0.f, 0.f);
                if (e < ne) {
                    const size_t nb = (size_t)sTgt[e] * DH + 4 * og;
                    const int4 as = *reinterpret_cast<const int4 *>(asel + nb);
                    const float4 go = *reinterpret_cast<const float4 *>(gout + nb);
                    const int asv[4] = {as.x, as.y, as.z, as.w};
                    const float gov[4] = {go.x, go.y, go.z, go.w};
                    const int src = sSrc[e];
                    float r[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int c = 4 * og + j;
                        const float h2 = elu(acc[i][j]);
                        const float hat = (h2 - sBn[c]) * sBn[DH + c];
                        const float dy = src == asv[j] ? gov[j] : 0.f;              // the edge the forward's max / min selected
                        const float dh2 = sBn[2 * DH + c] * sBn[DH + c] * (dy - sBn[3 * DH + c] - hat * sBn[4 * DH + c]);
                        r[j] = dh2 * elu_grad_from_out(h2);
                    }
                    dz = make_float4(r[0], r[1], r[2], r[3]);
                }
                *reinterpret_cast<float4 *>(sDZ + e * EB_LD_DZ + 4 * og) = dz;
            }
        }
        __syncthreads();
        {   // dW2[c][k] += dz2[e][c] h1[e][k]: thread = 4 channels x 3 k
            const int cg = tid & 15, kg = tid >> 4;
            for (int e = 0; e < DT; ++e) {
                const float4 d = *reinterpret_cast<const float4 *>(sDZ + e * EB_LD_DZ + 4 * cg);
                const float h0 = sH1[e * LD_H1 + 3 * kg], h1v = sH1[e * LD_H1 + 3 * kg + 1], h2v = sH1[e * LD_H1 + 3 * kg + 2];
                accW[0][0] = fmaf(d.x, h0, accW[0][0]); accW[0][1] = fmaf(d.x, h1v, accW[0][1]); accW[0][2] = fmaf(d.x, h2v, accW[0][2]);
                accW[1][0] = fmaf(d.y, h0, accW[1][0]); accW[1][1] = fmaf(d.y, h1v, accW[1][1]); accW[1][2] = fmaf(d.y, h2v, accW[1][2]);
                accW[2][0] = fmaf(d.z, h0, accW[2][0]); accW[2][1] = fmaf(d.z, h1v, accW[2][1]); accW[2][2] = fmaf(d.z, h2v, accW[2][2]);
                accW[3][0] = fmaf(d.w, h0, accW[3][0]); accW[3][1] = fmaf(d.w, h1v, accW[3][1]); accW[3][2] = fmaf(d.w, h2v, accW[3][2]);
            }
            if (tid < DH)
                for (int e = 0; e < DT; ++e) accb += sDZ[e * EB_LD_DZ + tid];
        }
        {   // dh1[e][k] = sum_c dz2[e][c] W2[c][k]; dz1 = dh1 * ELU'(h1): thread = 4 rows x 6 k
            float a2[4][6];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int q = 0; q < 6; ++q) a2[i][q] = 0.f;
#pragma unroll 4
            for (int c = 0; c < DH; ++c) {
                const float2 w0 = *reinterpret_cast<const float2 *>(sW2 + c * DM + 6 * og);
                const float2 w1 = *reinterpret_cast<const float2 *>(sW2 + c * DM + 6 * og + 2);
                const float2 w2v = *reinterpret_cast<const float2 *>(sW2 + c * DM + 6 * og + 4);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float d = sDZ[(ng + 32 * i) * EB_LD_DZ + c];
                    a2[i][0] = fmaf(d, w0.x, a2[i][0]); a2[i][1] = fmaf(d, w0.y, a2[i][1]);
                    a2[i][2] = fmaf(d, w1.x, a2[i][2]); a2[i][3] = fmaf(d, w1.y, a2[i][3]);
                    a2[i][4] = fmaf(d, w2v.x, a2[i][4]); a2[i][5] = fmaf(d, w2v.y, a2[i][5]);
                }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int e = ng + 32 * i;
                if (e < ne) {
                    float *dst = dz1 + (size_t)(e0 + e) * DM + 6 * og;
#pragma unroll
                    for (int q = 0; q < 6; q += 2)
                        *reinterpret_cast<float2 *>(dst + q) =
                            make_float2(a2[i][q] * elu_grad_from_out(sH1[e * LD_H1 + 6 * og + q]),
                                        a2[i][q + 1] * elu_grad_from_out(sH1[e * LD_H1 + 6 * og + q + 1]));
                }
            }
        }
    }
    float *dst = part + (size_t)blockIdx.x * EB_PART;
    {
        const int cg = tid & 15, kg = tid >> 4;
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) dst[(4 * cg + i) * DM + 3 * kg + j] = accW[i][j];
        if (tid < DH) dst[DH * DM + tid] = accb;
    }
}

// warp per node: d[A | B]_i = [ sum_{p in row i} dz1[p] | sum_{p in row i} dz1[mirror[p]] ]   (fixed order: deterministic)
__global__ void __launch_bounds__(256) k_drn_gather_ab(const float *__restrict__ dz1, const int *__restrict__ rowptr,
                                                       const int *__restrict__ mirror, int N, const int *__restrict__ n_dev,
                                                       float *__restrict__ dab) {
    pdl_sync();
    const int n = eff_n(N, n_dev), lane = threadIdx.x & 31;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    for (int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += warps) {
        const int a = rowptr[i], b = rowptr[i + 1];
        float sa[3] = {0.f, 0.f, 0.f}, sb[3] = {0.f, 0.f, 0.f};
        for (int p = a; p < b; ++p) {
            const float *ra = dz1 + (size_t)p * DM, *rb = dz1 + (size_t)mirror[p] * DM;
#pragma unroll
            for (int q = 0; q < 3; ++q) { sa[q] += ra[lane + 32 * q]; sb[q] += rb[lane + 32 * q]; }
        }
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            dab[(size_t)i * AB_W + lane + 32 * q] = sa[q];
            dab[(size_t)i * AB_W + DM + lane + 32 * q] = sb[q];
        }
    }
}

// node side of an EdgeConv stage: dX = d[A|B] . [W1a - W1b | W1b] (per 128-node tile) and the per-CTA partials
// G = d[A|B]^T X  [192][64], db1 = sum_n dA.   dW1[o][k<64] = G[o][k], dW1[o][64+k] = G[96+o][k] - G[o][k].
constexpr int NB_LD_D = 196;
constexpr size_t NB_SMEM = (size_t)(AB_W * DH + DT * NB_LD_D + DT * LD_B) * 4;
constexpr int NB_PART = AB_W * DH + DM;
__global__ void __launch_bounds__(DTHREADS, 1) k_drn_node_bwd(const float *__restrict__ dab, const float *__restrict__ x, int N,
                                                              const int *__restrict__ n_dev, const float *__restrict__ cw1,
                                                              float *__restrict__ dx, float *__restrict__ part) {
    extern __shared__ __align__(16) float s_nb[];
    float *sWo = s_nb, *sD = sWo + AB_W * DH, *sX = sD + DT * NB_LD_D;
    const int tid = threadIdx.x;
    for (int t = tid; t < DM * DH; t += DTHREADS) {       // sWo[o][k]: row o of [W1a - W1b | W1b] (what o contributes to x_k)
        const int o = t / DH, k = t - o * DH;
        const float wa = cw1[(size_t)o * 2 * DH + k], wb = cw1[(size_t)o * 2 * DH + DH + k];
        sWo[o * DH + k] = wa - wb;
        sWo[(DM + o) * DH + k] = wb;
    }
    pdl_sync();
    const int n = eff_n(N, n_dev);
    const int ng = tid >> 4, og = tid & 15;
    const int og2 = tid & 31, kg = tid >> 5;              // G block: o = 6 og2 .. + 5, k = 4 kg .. + 3
    float accG[6][4], accb = 0.f;
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) accG[i][j] = 0.f;
    for (int n0 = blockIdx.x * DT; n0 < n; n0 += gridDim.x * DT) {
        const int rows = min(DT, n - n0);
        __syncthreads();
        for (int t = tid; t < DT * (AB_W / 4); t += DTHREADS) {
            const int r = t / (AB_W / 4), q = t - r * (AB_W / 4);
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (r < rows) v = reinterpret_cast<const float4 *>(dab + (size_t)(n0 + r) * AB_W)[q];
            *reinterpret_cast<float4 *>(sD + r * NB_LD_D + 4 * q) = v;
        }
        for (int t = tid; t < DT * (DH / 4); t += DTHREADS) {
            const int r = t >> 4, q = t & 15;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (r < rows) v = reinterpret_cast<const float4 *>(x + (size_t)(n0 + r) * DH)[q];
            *reinterpret_cast<float4 *>(sX + r * LD_B + 4 * q) = v;
        }
        __syncthreads();
        {
            float acc[4][4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
            tile_gemm<AB_W, DH, NB_LD_D>(sD, sWo, ng, og, acc);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int r = ng + 32 * i;
                if (r < rows)
                    *reinterpret_cast<float4 *>(dx + (size_t)(n0 + r) * DH + 4 * og) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
            }
        }
        for (int r = 0; r < DT; ++r) {
            const float2 d0 = *reinterpret_cast<const float2 *>(sD + r * NB_LD_D + 6 * og2);
            const float2 d1 = *reinterpret_cast<const float2 *>(sD + r * NB_LD_D + 6 * og2 + 2);
            const float2 d2 = *reinterpret_cast<const float2 *>(sD + r * NB_LD_D + 6 * og2 + 4);
            const float4 xv = *reinterpret_cast<const float4 *>(sX + r * LD_B + 4 * kg);
            const float dv[6] = {d0.x, d0.y, d1.x, d1.y, d2.x, d2.y};
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                accG[i][0] = fmaf(dv[i], xv.x, accG[i][0]); accG[i][1] = fmaf(dv[i], xv.y, accG[i][1]);
                accG[i][2] = fmaf(dv[i], xv.z, accG[i][2]); accG[i][3] = fmaf(dv[i], xv.w, accG[i][3]);
            }
        }
        if (tid < DM)
            for (int r = 0; r < DT; ++r) accb += sD[r * NB_LD_D + tid];
    }
    float *dst = part + (size_t)blockIdx.x * NB_PART;
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) dst[(6 * og2 + i) * DH + 4 * kg + j] = accG[i][j];
    if (tid < DM) dst[AB_W * DH + tid] = accb;
}

// max_pool backward: the pooled gradient of a cluster goes, per channel, to its larger member (the lower index on a tie)
__global__ void __launch_bounds__(256) k_drn_pool_bwd(const float *__restrict__ x1, const int *__restrict__ cluster,
                                                      const int *__restrict__ partner, const int *__restrict__ scan, int N,
                                                      const float *__restrict__ dx1p, float *__restrict__ dx1) {
    pdl_sync();
    const long long total = (long long)N * DH;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        const int u = (int)(t >> 6), c = (int)(t & (DH - 1));
        const int cu = cluster[u], v = partner[u];
        const float g = dx1p[(size_t)scan[cu] * DH + c];
        float out = g;
        if (v >= 0) {
            const int lo = min(u, v), hi = max(u, v);
            const bool lo_wins = x1[(size_t)lo * DH + c] >= x1[(size_t)hi * DH + c];
            out = (u == lo) == lo_wins ? g : 0.f;
        }
        dx1[t] = out;
    }
}

// inputnet backward: recompute the three layers of a 128-node tile (k_drn_node's code), back-propagate d x0
constexpr int IB_PART = DH * DH + DH + DH * 32 + DH + 32 * DI + 32 + 16;    // dW3 | db3 | dW2 | db2 | dW1 | db1 | dnorm
constexpr size_t IB_SMEM = (size_t)(NS_W1 + 32 + NS_W2 + DH + NS_W3 + DH + 16 + DH * DH + DH * 32 + NS_IN + NS_A + NS_B + NS_C +
                                    DT * 12) * 4;
__global__ void __launch_bounds__(DTHREADS, 1) k_drn_input_bwd(const float *__restrict__ x, int x_stride, int N, NodeParams p,
                                                               const float *__restrict__ dx0, float *__restrict__ part) {
    extern __shared__ __align__(16) float s_ib[];
    float *sW1 = s_ib, *sb1 = sW1 + NS_W1, *sW2 = sb1 + 32, *sb2 = sW2 + NS_W2, *sW3 = sb2 + DH, *sb3 = sW3 + NS_W3,
          *sNorm = sb3 + DH, *sW3r = sNorm + 16, *sW2r = sW3r + DH * DH, *sIn = sW2r + DH * 32, *sA = sIn + NS_IN,
          *sB = sA + NS_A, *sC = sB + NS_B, *sXn = sC + NS_C;
    const int tid = threadIdx.x;
    stage_kmajor(p.w1, 32, DI, DI, sW1, 32, 0);
    stage_kmajor(p.w2, DH, 32, 32, sW2, DH, 0);
    stage_kmajor(p.w3, DH, DH, DH, sW3, DH, 0);
    for (int t = tid; t < DH * DH; t += DTHREADS) sW3r[t] = p.w3[t];        // row-major [o][k]: da2 = dz3 . W3
    for (int t = tid; t < DH * 32; t += DTHREADS) sW2r[t] = p.w2[t];        // row-major [o][k], k < 32: da1 = dz2 . W2
    if (tid < 32) sb1[tid] = p.b1[tid];
    if (tid < DH) { sb2[tid] = p.b2[tid]; sb3[tid] = p.b3[tid]; }
    if (tid < DI) sNorm[tid] = p.norm[tid];
    pdl_sync();
    const int ng = tid >> 4, og = tid & 15;
    float aW3[4][2], aW2[4], ab3 = 0.f, ab2 = 0.f, aW1 = 0.f, ab1 = 0.f, anorm = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) { aW3[i][0] = aW3[i][1] = 0.f; aW2[i] = 0.f; }
    for (int n0 = blockIdx.x * DT; n0 < N; n0 += gridDim.x * DT) {
        const int rows = min(DT, N - n0);
        __syncthreads();
        for (int t = tid; t < DT * DI; t += DTHREADS) {
            const int r = t / DI, k = t - r * DI;
            sIn[r * LD_IN + k] = r < rows ? x[(size_t)(n0 + r) * x_stride + k] * sNorm[k] : 0.f;
        }
        __syncthreads();
        {   // forward, exactly as k_drn_node<true>
            const int node = tid & (DT - 1), o0 = (tid >> 7) * 8;
            float acc[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] = sb1[o0 + j];
#pragma unroll
            for (int k = 0; k < DI; ++k) {
                const float v = sIn[node * LD_IN + k];
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] = fmaf(v, sW1[k * 32 + o0 + j], acc[j]);
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) sA[node * LD_A + o0 + j] = elu(acc[j]);
        }
        __syncthreads();
        {
            float acc[4][4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = sb2[4 * og + j];
            tile_gemm<32, DH, LD_A>(sA, sW2, ng, og, acc);
#pragma unroll
            for (int i = 0; i < 4; ++i)
                *reinterpret_cast<float4 *>(sB + (ng + 32 * i) * LD_B + 4 * og) =
                    make_float4(elu(acc[i][0]), elu(acc[i][1]), elu(acc[i][2]), elu(acc[i][3]));
        }
        __syncthreads();
        {   // layer 3 forward and dz3 = d x0 * ELU'(x0) in one go (sC holds dz3)
            float acc[4][4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = sb3[4 * og + j];
            tile_gemm<DH, DH, LD_B>(sB, sW3, ng, og, acc);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int r = ng + 32 * i;
                float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
                if (r < rows) g = *reinterpret_cast<const float4 *>(dx0 + (size_t)(n0 + r) * DH + 4 * og);
                *reinterpret_cast<float4 *>(sC + r * LD_B + 4 * og) =
                    make_float4(g.x * elu_grad_from_out(elu(acc[i][0])), g.y * elu_grad_from_out(elu(acc[i][1])),
                                g.z * elu_grad_from_out(elu(acc[i][2])), g.w * elu_grad_from_out(elu(acc[i][3])));
            }
        }
        __syncthreads();
        {   // dW3[o][k] += dz3[n][o] a2[n][k]: thread = 4 o x 2 k;  db3
            const int kg = tid >> 4;
            for (int r = 0; r < DT; ++r) {
                const float4 d = *reinterpret_cast<const float4 *>(sC + r * LD_B + 4 * og);
                const float2 a = *reinterpret_cast<const float2 *>(sB + r * LD_B + 2 * kg);
                aW3[0][0] = fmaf(d.x, a.x, aW3[0][0]); aW3[0][1] = fmaf(d.x, a.y, aW3[0][1]);
                aW3[1][0] = fmaf(d.y, a.x, aW3[1][0]); aW3[1][1] = fmaf(d.y, a.y, aW3[1][1]);
                aW3[2][0] = fmaf(d.z, a.x, aW3[2][0]); aW3[2][1] = fmaf(d.z, a.y, aW3[2][1]);
                aW3[3][0] = fmaf(d.w, a.x, aW3[3][0]); aW3[3][1] = fmaf(d.w, a.y, aW3[3][1]);
            }
            if (tid < DH)
                for (int r = 0; r < DT; ++r) ab3 += sC[r * LD_B + tid];
        }
        float da2[4][4];
        {   // da2 = dz3 . W3 (reduction over the output index o: W3 row-major is "K-major" here)
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) da2[i][j] = 0.f;
            tile_gemm<DH, DH, LD_B>(sC, sW3r, ng, og, da2);
        }
        __syncthreads();                                   // everyone is done with dz3 (sC) and a2 is still needed below
#pragma unroll
        for (int i = 0; i < 4; ++i) {                      // dz2 = da2 * ELU'(a2) -> sC
            const int r = ng + 32 * i;
            const float4 a = *reinterpret_cast<const float4 *>(sB + r * LD_B + 4 * og);
            *reinterpret_cast<float4 *>(sC + r * LD_B + 4 * og) =
                make_float4(da2[i][0] * elu_grad_from_out(a.x), da2[i][1] * elu_grad_from_out(a.y),
                            da2[i][2] * elu_grad_from_out(a.z), da2[i][3] * elu_grad_from_out(a.w));
        }
        __syncthreads();
        {   // dW2[o][k] += dz2[n][o] a1[n][k]: thread = 4 o x 1 k (k < 32);  db2
            const int k = tid >> 4;
            for (int r = 0; r < DT; ++r) {
                const float4 d = *reinterpret_cast<const float4 *>(sC + r * LD_B + 4 * og);
                const float a = sA[r * LD_A + k];
                aW2[0] = fmaf(d.x, a, aW2[0]); aW2[1] = fmaf(d.y, a, aW2[1]);
                aW2[2] = fmaf(d.z, a, aW2[2]); aW2[3] = fmaf(d.w, a, aW2[3]);
            }
            if (tid < DH)
                for (int r = 0; r < DT; ++r) ab2 += sC[r * LD_B + tid];
        }
        float da1[4][4];
        const int ng8 = tid >> 3, og8 = tid & 7;           // 32 outputs: threads 0..255 own 4 rows x 4 k
        if (tid < 256) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) da1[i][j] = 0.f;
            tile_gemm<DH, 32, LD_B>(sC, sW2r, ng8, og8, da1);
        }
        __syncthreads();                                   // a1 (sA) consumed by dW2: overwrite it with dz1
        if (tid < 256) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int r = ng8 + 32 * i;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float a = sA[r * LD_A + 4 * og8 + j];
                    sA[r * LD_A + 4 * og8 + j] = da1[i][j] * elu_grad_from_out(a);
                }
            }
        }
        __syncthreads();
        if (tid < 32 * DI) {                               // dW1[o][k] += dz1[n][o] xn[n][k]
            const int o = tid / DI, k = tid - o * DI;
            for (int r = 0; r < DT; ++r) aW1 = fmaf(sA[r * LD_A + o], sIn[r * LD_IN + k], aW1);
        } else if (tid >= 352 && tid < 384) {              // db1
            const int o = tid - 352;
            for (int r = 0; r < DT; ++r) ab1 += sA[r * LD_A + o];
        }
        if (tid < DT) {                                    // d xn[n][k] = sum_o dz1[n][o] W1[o][k]; d norm[k] += d xn * x
            const int r = tid;
#pragma unroll
            for (int k = 0; k < DI; ++k) {
                float acc = 0.f;
                for (int o = 0; o < 32; ++o) acc = fmaf(sA[r * LD_A + o], sW1[k * 32 + o], acc);
                sXn[r * 12 + k] = r < rows ? acc * x[(size_t)(n0 + r) * x_stride + k] : 0.f;
            }
        }
        __syncthreads();
        if (tid < DI)
            for (int r = 0; r < DT; ++r) anorm += sXn[r * 12 + tid];
    }
    float *dst = part + (size_t)blockIdx.x * IB_PART;
    {
        const int kg = tid >> 4;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            dst[(4 * og + i) * DH + 2 * kg] = aW3[i][0];
            dst[(4 * og + i) * DH + 2 * kg + 1] = aW3[i][1];
            dst[DH * DH + DH + (4 * og + i) * 32 + kg] = aW2[i];
        }
        if (tid < DH) { dst[DH * DH + tid] = ab3; dst[DH * DH + DH + DH * 32 + tid] = ab2; }
        const int o1 = DH * DH + DH + DH * 32 + DH;
        if (tid < 32 * DI) dst[o1 + tid] = aW1;
        else if (tid >= 352 && tid < 384) dst[o1 + 32 * DI + tid - 352] = ab1;
        if (tid < 16) dst[o1 + 32 * DI + 32 + tid] = tid < DI ? anorm : 0.f;
    }
}

// summed partials -> parameter gradients
__global__ void __launch_bounds__(256) k_drn_fin_conv_w1(const float *__restrict__ sum, float *__restrict__ dw1, float *__restrict__ db1) {
    pdl_sync();
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < DM * 2 * DH + DM; t += gridDim.x * blockDim.x) {
        if (t < DM * 2 * DH) {
            const int o = t / (2 * DH), k = t - o * 2 * DH;
            dw1[t] = k < DH ? sum[o * DH + k] : sum[(DM + o) * DH + k - DH] - sum[o * DH + k - DH];
        } else {
            db1[t - DM * 2 * DH] = sum[AB_W * DH + t - DM * 2 * DH];
        }
    }
}
__global__ void __launch_bounds__(256) k_drn_copy_seg(const float *__restrict__ src, float *__restrict__ dst, int n) {
    pdl_sync();
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) dst[t] = src[t];
}

}  // namespace
}  // namespace gm

using namespace gm;

extern "C" {

size_t gm_drn_ws_bytes(int N, int B, int k) { return drn_layout(N < 0 ? 0 : N, B < 0 ? 0 : B, k < 1 ? 1 : k, nullptr, nullptr); }

int gm_drn_forward(const GmDrnParams *p, const float *x, int x_stride, const int *ptr, int N, int B,
                   int max_event_nodes, int training, int precision, float momentum, float eps, void *ws,
                   size_t ws_bytes, float *out, float *net_out, GmDrnView *view_out, void *stream) {
    if (!p || !ws || N < 1 || B < 1 || !x || !ptr || !out || x_stride < DI) return GM_ERR_INVALID_ARGUMENT;
    if (p->k < 1 || p->k + 1 > 32 || p->output_dim < 1 || p->output_dim > 8) return GM_ERR_UNSUPPORTED;
    if (precision != GM_PREC_FP32 && precision != GM_PREC_BF16) return GM_ERR_UNSUPPORTED;   // tf32: not built for the DRN
    if (net_out && p->output_dim != 2) return GM_ERR_INVALID_ARGUMENT;
    if (ws_bytes < drn_layout(N, B, p->k, nullptr, nullptr)) return GM_ERR_CAPACITY;
    cudaStream_t st = as_stream(stream);
    DrnWs w;
    drn_layout(N, B, p->k, (char *)ws, &w);
    const int K = p->k + 1;
    static bool attr = false;
    if (!attr) {
        GM_CUDA_TRY(cudaFuncSetAttribute(k_drn_node<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)NODE_SMEM_FULL));
        GM_CUDA_TRY(cudaFuncSetAttribute(k_drn_node<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)NODE_SMEM_BASE));
        attr = true;
    }
    const int node_grid = persistent_grid(((long long)N + DT - 1) / DT, 1);
    // ---- stage 1 (:82-92)
    NodeParams np{p->datanorm, p->in_w[0], p->in_b[0], p->in_w[1], p->in_b[1], p->in_w[2], p->in_b[2], p->conv[0].w1, p->conv[0].b1};
    {
        TimedRegion tr(GM_TIMER_DRN_NODE, st);
        GM_LAUNCH_PDL(k_drn_node<true>, node_grid, DTHREADS, NODE_SMEM_FULL, st, x, x_stride, N, (const int *)nullptr, np, w.x0, w.ab);
        GM_LAUNCH_CHECK();
    }
    {
        TimedRegion tr(GM_TIMER_DRN_KNN, st);
        GM_TRY(knn_search_table(w.x0, DH, DH, ptr, N, B, K, 0, w.nbr, w.deg, nullptr, st));
    }
    {
        TimedRegion tr(GM_TIMER_DRN_GRAPH, st);
        GM_TRY(drn_to_undirected(nullptr, nullptr, w.nbr, w.deg, K, N, nullptr, w.und, w.und_bytes, w.rowptr1, w.col1, w.row1,
                                 w.ecap, nullptr, st));
    }
    GM_TRY(edge_stage(w, p->conv[0], w.rowptr1, w.col1, w.row1, N, nullptr, training, precision, momentum, eps, w.bnss, w.x1, w.vsel1, w.asel1, st));
    {
        TimedRegion tr(GM_TIMER_DRN_POOL, st);
        GM_TRY(drn_normalized_cut(w.x1, DH, DH, w.rowptr1, w.col1, w.row1, N, w.ecap, w.weight1, st));
        GM_TRY(drn_graclus(w.rowptr1, w.col1, w.weight1, ptr, N, B, max_event_nodes, nullptr, w.cluster1, w.partner1, st));
        GM_TRY(drn_max_pool(w.x1, DH, w.cluster1, w.partner1, ptr, N, B, nullptr, w.scan, w.scan_tiles, w.x1p, w.ptr1p, nullptr, st));
    }
    // ---- stage 2 (:95-96) on the M = ptr1p[B] pooled nodes; M never leaves the device
    const int *m_dev = w.ptr1p + B;
    NodeParams np2{nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, p->conv[1].w1, p->conv[1].b1};
    {
        TimedRegion tr(GM_TIMER_DRN_NODE, st);
        GM_LAUNCH_PDL(k_drn_node<false>, node_grid, DTHREADS, NODE_SMEM_BASE, st, (const float *)w.x1p, DH, N, m_dev, np2,
                      (float *)nullptr, w.ab);
        GM_LAUNCH_CHECK();
    }
    {
        TimedRegion tr(GM_TIMER_DRN_KNN, st);
        GM_TRY(knn_search_table(w.x1p, DH, DH, w.ptr1p, N, B, K, 0, w.nbr, w.deg, m_dev, st));
    }
    {
        TimedRegion tr(GM_TIMER_DRN_GRAPH, st);
        GM_TRY(drn_to_undirected(nullptr, nullptr, w.nbr, w.deg, K, N, m_dev, w.und, w.und_bytes, w.rowptr2, w.col2, w.row2,
                                 w.ecap, nullptr, st));
    }
    GM_TRY(edge_stage(w, p->conv[1], w.rowptr2, w.col2, w.row2, N, m_dev, training, precision, momentum, eps, w.bnss + 4 * DH, w.x2, w.vsel2, w.asel2, st));
    // ---- :98-104: the second clustering cannot change a global maximum; pool per event and run the output MLP
    HeadParams hp{p->out_w[0], p->out_b[0], p->out_w[1], p->out_b[1], p->out_w[2], p->out_b[2], p->output_dim};
    {
        TimedRegion tr(GM_TIMER_DRN_POOL, st);
        GM_LAUNCH_PDL(k_drn_head, B, HEAD_PARTS * DH, 0, st, (const float *)w.x2, (const int *)w.ptr1p, B, hp, w.pooled, out, net_out);
        GM_LAUNCH_CHECK();
    }
    if (view_out) {
        GmDrnView v{};
        v.x0 = w.x0; v.rowptr1 = w.rowptr1; v.col1 = w.col1; v.row1 = w.row1; v.x1 = w.x1; v.weight1 = w.weight1;
        v.cluster1 = w.cluster1; v.partner1 = w.partner1; v.x1p = w.x1p; v.ptr1p = w.ptr1p; v.rowptr2 = w.rowptr2;
        v.col2 = w.col2; v.row2 = w.row2; v.x2 = w.x2; v.pooled = w.pooled; v.edge_capacity = w.ecap;
        v.asel1 = w.asel1; v.asel2 = w.asel2;
        *view_out = v;
    }
    return GM_OK;
}

}  // extern "C"

namespace gm {
namespace {

// The wrapper of model/net.py:31-36 fused with a regression loss on the MET vector it parametrises:
//   met = softplus(out0), phi = pi (2 sigmoid(out1) - 1), loss = 0.5 mean_e[(met cos phi - t0)^2 + (met sin phi - t1)^2]
// (the reference keeps no loss for this variant; this is the Cartesian MSE against genMET px, py = truth columns 0, 1).
// One CTA: B events, fixed-order block reduction; also d loss / d out [B,2] so the backward needs no extra pass.
__global__ void __launch_bounds__(256) k_drn_met_loss(const float *__restrict__ out, const float *__restrict__ truth,
                                                      int truth_stride, int B, float grad_scale, float *__restrict__ net_out,
                                                      float *__restrict__ loss_out, float *__restrict__ dout) {
    __shared__ double s_red[256];
    pdl_sync();
    double acc = 0.0;
    for (int e = threadIdx.x; e < B; e += 256) {
        const float o0 = out[2 * e], o1 = out[2 * e + 1];
        const float met = o0 > 20.f ? o0 : log1pf(expf(o0));
        const float sg = sigmoidf(o1), phi = 3.14159265358979323846f * (2.f * sg - 1.f);
        float sp, cp;
        sincosf(phi, &sp, &cp);
        const float rx = met * cp - truth[(size_t)e * truth_stride], ry = met * sp - truth[(size_t)e * truth_stride + 1];
        acc += 0.5 * ((double)rx * rx + (double)ry * ry);
        if (net_out) { net_out[2 * e] = met; net_out[2 * e + 1] = phi; }
        if (dout) {
            const float dmet = (rx * cp + ry * sp) * grad_scale, dphi = met * (ry * cp - rx * sp) * grad_scale;
            dout[2 * e] = dmet * sigmoidf(o0);                                   // d softplus
            dout[2 * e + 1] = dphi * 2.f * 3.14159265358979323846f * sg * (1.f - sg);
        }
    }
    s_red[threadIdx.x] = acc;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) s_red[threadIdx.x] += s_red[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) loss_out[0] = (float)(s_red[0] / (double)B);
}

// ------------------------------------------------------------------------------------------ backward: host side
struct DrnBwdWs {
    float *dz1, *dab, *gA, *gB, *sums, *part, *psum;
    int *mirror;
    size_t part_floats;
};
constexpr int BWD_MAX_CTAS = 148;
size_t drn_bwd_layout(int N, int B, int k, char *base, DrnBwdWs *w) {
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return base ? base + o : (char *)nullptr; };
    const size_t n = (size_t)(N > 0 ? N : 1), K = (size_t)k + 1, ecap = 2 * n * K;
    DrnBwdWs t{};
    t.dz1 = (float *)take(ecap * DM * 4);
    t.mirror = (int *)take(ecap * 4);
    t.dab = (float *)take(n * AB_W * 4);
    t.gA = (float *)take(n * DH * 4);
    t.gB = (float *)take(n * DH * 4);
    t.sums = (float *)take(2 * DH * 4);
    size_t pf = (size_t)BWD_MAX_CTAS * NB_PART;
    if ((size_t)BWD_MAX_CTAS * EB_PART > pf) pf = (size_t)BWD_MAX_CTAS * EB_PART;
    if ((size_t)BWD_MAX_CTAS * IB_PART > pf) pf = (size_t)BWD_MAX_CTAS * IB_PART;
    if ((size_t)(B > 0 ? B : 1) * HB_LEN > pf) pf = (size_t)(B > 0 ? B : 1) * HB_LEN;
    t.part_floats = pf;
    t.part = (float *)take(pf * 4);
    t.psum = (float *)take((size_t)NB_PART * 4);
    if (w) *w = t;
    return off;
}

int sum_rows(const float *src, int rows, int len, float *dst, cudaStream_t st) {
    GM_LAUNCH_PDL(k_sum_rows, (len + 255) / 256, 256, 0, st, src, rows, len, dst);
    GM_LAUNCH_CHECK();
    return GM_OK;
}
int copy_seg(const float *src, float *dst, int n, cudaStream_t st) {
    if (!dst) return GM_OK;
    GM_LAUNCH_PDL(k_drn_copy_seg, (n + 255) / 256, 256, 0, st, src, dst, n);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

// one EdgeConv stage backwards: gout = d x_out [n,64] -> gin = d x_in [n,64]; conv gradients written
int edge_stage_bwd(const DrnWs &f, const DrnBwdWs &w, const GmDrnConv &cv, const GmDrnConv &gcv, const float *x_in,
                   const float *vsel, const int *asel, const float *bnss, const int *rowptr, const int *col, const int *row, int N,
                   const int *n_dev, const float *gout, float *gin, cudaStream_t st) {
    const int node_grid = persistent_grid(((long long)N + DT - 1) / DT, 1);
    {
    TimedRegion tr0(GM_TIMER_DRN_BWD, st);
    const int sgrid = persistent_grid(((long long)N + BS_PARTS - 1) / BS_PARTS, 1);
    GM_LAUNCH_PDL(k_drn_bn_sums, sgrid, BS_PARTS * DH, 0, st, gout, vsel, rowptr, N, n_dev, bnss, (double *)w.part);
    GM_LAUNCH_CHECK();
    GM_LAUNCH_PDL(k_drn_bn_sums_fin, 1, 2 * DH, 0, st, (const double *)w.part, sgrid, w.sums, gcv.bn_weight, gcv.bn_bias);
    GM_LAUNCH_CHECK();
    // [A | B] of this stage again (the forward keeps one buffer for both stages)
    NodeParams np{nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, cv.w1, cv.b1};
    GM_LAUNCH_PDL(k_drn_node<false>, node_grid, DTHREADS, NODE_SMEM_BASE, st, x_in, DH, N, n_dev, np, (float *)nullptr, f.ab);
    GM_LAUNCH_CHECK();
    GM_LAUNCH_PDL(k_drn_mirror, persistent_grid((f.ecap + 255) / 256, 8), 256, 0, st, rowptr, col, row, N, f.ecap, w.mirror);
    GM_LAUNCH_CHECK();
    }
    const int eg = persistent_grid(f.tiles_cap, 1);
    {
    TimedRegion tre(GM_TIMER_DRN_EDGE_BWD, st);
    GM_LAUNCH_PDL(k_drn_edge_bwd, eg, DTHREADS, EB_SMEM, st, (const float *)f.ab, rowptr, row, col, N, f.ecap,
                  (const float *)cv.w2, (const float *)cv.b2, bnss, (const float *)w.sums, asel, gout, w.dz1, w.part);
    GM_LAUNCH_CHECK();
    }
    TimedRegion trr(GM_TIMER_DRN_BWD, st);
    GM_TRY(sum_rows(w.part, eg, EB_PART, w.psum, st));
    GM_TRY(copy_seg(w.psum, gcv.w2, DH * DM, st));
    GM_TRY(copy_seg(w.psum + DH * DM, gcv.b2, DH, st));
    GM_LAUNCH_PDL(k_drn_gather_ab, persistent_grid(((long long)N * 32 + 255) / 256, 8), 256, 0, st, (const float *)w.dz1, rowptr,
                  (const int *)w.mirror, N, n_dev, w.dab);
    GM_LAUNCH_CHECK();
    GM_LAUNCH_PDL(k_drn_node_bwd, node_grid, DTHREADS, NB_SMEM, st, (const float *)w.dab, x_in, N, n_dev, (const float *)cv.w1,
                  gin, w.part);
    GM_LAUNCH_CHECK();
    GM_TRY(sum_rows(w.part, node_grid, NB_PART, w.psum, st));
    if (gcv.w1 && gcv.b1) {
        GM_LAUNCH_PDL(k_drn_fin_conv_w1, (DM * 2 * DH + DM + 255) / 256, 256, 0, st, (const float *)w.psum, gcv.w1, gcv.b1);
        GM_LAUNCH_CHECK();
    }
    return GM_OK;
}

}  // namespace
}  // namespace gm

using namespace gm;

extern "C" {

size_t gm_drn_bwd_ws_bytes(int N, int B, int k) { return drn_bwd_layout(N < 0 ? 0 : N, B < 0 ? 0 : B, k < 1 ? 1 : k, nullptr, nullptr); }

int gm_drn_backward(const GmDrnParams *p, const GmDrnParams *grads, const float *x, int x_stride, const int *ptr, int N,
                    int B, const void *fwd_ws, size_t fwd_ws_bytes, void *bwd_ws, size_t bwd_ws_bytes,
                    const float *grad_out, void *stream) {
    if (!p || !grads || !fwd_ws || !bwd_ws || N < 1 || B < 1 || !x || !ptr || !grad_out || x_stride < DI) return GM_ERR_INVALID_ARGUMENT;
    if (p->k < 1 || p->k + 1 > 32 || p->output_dim < 1 || p->output_dim > 8) return GM_ERR_UNSUPPORTED;
    if (fwd_ws_bytes < drn_layout(N, B, p->k, nullptr, nullptr) || bwd_ws_bytes < drn_bwd_layout(N, B, p->k, nullptr, nullptr))
        return GM_ERR_CAPACITY;
    cudaStream_t st = as_stream(stream);
    DrnWs f;
    drn_layout(N, B, p->k, (char *)fwd_ws, &f);
    DrnBwdWs w;
    drn_bwd_layout(N, B, p->k, (char *)bwd_ws, &w);
    static bool attr = false;
    if (!attr) {
        GM_CUDA_TRY(cudaFuncSetAttribute(k_drn_edge_bwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)EB_SMEM));
        GM_CUDA_TRY(cudaFuncSetAttribute(k_drn_node_bwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)NB_SMEM));
        GM_CUDA_TRY(cudaFuncSetAttribute(k_drn_input_bwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)IB_SMEM));
        GM_CUDA_TRY(cudaFuncSetAttribute(k_drn_node<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)NODE_SMEM_BASE));
        attr = true;
    }
    const int *m_dev = f.ptr1p + B;
    // ---- output MLP + global max pool
    HeadParams hp{p->out_w[0], p->out_b[0], p->out_w[1], p->out_b[1], p->out_w[2], p->out_b[2], p->output_dim};
    GM_LAUNCH_PDL(k_drn_head_bwd, B, 256, 0, st, (const float *)f.x2, (const int *)f.ptr1p, B, (const float *)f.pooled, hp,
                  grad_out, w.part, w.gA);
    GM_LAUNCH_CHECK();
    GM_TRY(sum_rows(w.part, B, HB_LEN, w.psum, st));
    {
        const int od = p->output_dim, o2 = DH * DH + DH, o3 = o2 + (DH / 2) * DH + DH / 2;
        GM_TRY(copy_seg(w.psum, grads->out_w[0], DH * DH, st));
        GM_TRY(copy_seg(w.psum + DH * DH, grads->out_b[0], DH, st));
        GM_TRY(copy_seg(w.psum + o2, grads->out_w[1], (DH / 2) * DH, st));
        GM_TRY(copy_seg(w.psum + o2 + (DH / 2) * DH, grads->out_b[1], DH / 2, st));
        GM_TRY(copy_seg(w.psum + o3, grads->out_w[2], od * (DH / 2), st));
        GM_TRY(copy_seg(w.psum + o3 + 8 * (DH / 2), grads->out_b[2], od, st));
    }
    // ---- stage 2 (M pooled nodes), max_pool, stage 1
    GM_TRY(edge_stage_bwd(f, w, p->conv[1], grads->conv[1], f.x1p, f.vsel2, f.asel2, f.bnss + 4 * DH, f.rowptr2, f.col2, f.row2, N, m_dev,
                          w.gA, w.gB, st));
    GM_LAUNCH_PDL(k_drn_pool_bwd, persistent_grid(((long long)N * DH + 255) / 256, 8), 256, 0, st, (const float *)f.x1,
                  (const int *)f.cluster1, (const int *)f.partner1, (const int *)f.scan, N, (const float *)w.gB, w.gA);
    GM_LAUNCH_CHECK();
    GM_TRY(edge_stage_bwd(f, w, p->conv[0], grads->conv[0], f.x0, f.vsel1, f.asel1, f.bnss, f.rowptr1, f.col1, f.row1, N, nullptr, w.gA,
                          w.gB, st));
    // ---- inputnet + datanorm
    NodeParams np{p->datanorm, p->in_w[0], p->in_b[0], p->in_w[1], p->in_b[1], p->in_w[2], p->in_b[2], nullptr, nullptr};
    const int ig = persistent_grid(((long long)N + DT - 1) / DT, 1);
    GM_LAUNCH_PDL(k_drn_input_bwd, ig, DTHREADS, IB_SMEM, st, x, x_stride, N, np, (const float *)w.gB, w.part);
    GM_LAUNCH_CHECK();
    GM_TRY(sum_rows(w.part, ig, IB_PART, w.psum, st));
    {
        const int o2 = DH * DH + DH, o1 = o2 + DH * 32 + DH;
        GM_TRY(copy_seg(w.psum, grads->in_w[2], DH * DH, st));
        GM_TRY(copy_seg(w.psum + DH * DH, grads->in_b[2], DH, st));
        GM_TRY(copy_seg(w.psum + o2, grads->in_w[1], DH * 32, st));
        GM_TRY(copy_seg(w.psum + o2 + DH * 32, grads->in_b[1], DH, st));
        GM_TRY(copy_seg(w.psum + o1, grads->in_w[0], 32 * DI, st));
        GM_TRY(copy_seg(w.psum + o1 + 32 * DI, grads->in_b[0], 32, st));
        GM_TRY(copy_seg(w.psum + o1 + 32 * DI + 32, grads->datanorm, DI, st));
    }
    return GM_OK;
}

int gm_drn_met_loss(const float *out, const float *truth, int truth_stride, int B, float *net_out, float *loss_out,
                    float *grad_out, void *stream) {
    if (B < 1 || !out || !truth || !loss_out || truth_stride < 2) return GM_ERR_INVALID_ARGUMENT;
    GM_LAUNCH_PDL(k_drn_met_loss, 1, 256, 0, as_stream(stream), out, truth, truth_stride, B, 1.f / (float)B, net_out, loss_out,
                  grad_out);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

}  // extern "C"
