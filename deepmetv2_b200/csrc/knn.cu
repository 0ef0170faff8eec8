// k-nearest-neighbour graph per event in a D-dimensional (latent) space -- SURVEY.md §8 rows f4 / f1:
//   knn_graph(emb, k=20, batch, loop=True)         /root/reference/model/graph_met_network.py:63 (sketched there)
//   knn_graph(x, self.k, batch, loop=False, flow)  /root/reference/model/dynamic_reduction_network.py:86,95
// [upstream torch_cluster.knn_graph] semantics restated in oracle/knn_bruteforce.c; this file is its CUDA form and is
// bit-identical to it: dist = sum_d (x_j[d] - x_i[d])^2 accumulated in fp32 in dimension order as fmaf(df, df, dist)
// (what nvcc makes of upstream's `dist += df * df`), candidates scanned in ascending index, a candidate enters the list only when strictly
// closer than the current worst entry => ties go to the lower index; the K best (K = k + !loop) come out ordered by
// (distance, index), the self pair is dropped afterwards when loop == 0 (so a node with > k exact duplicates at lower
// indices keeps k + 1 neighbours, as upstream's `row != col` mask does).
//
// Layout: one warp owns QPW queries whose coordinates sit in registers; the CTA streams the event's candidates
// through shared memory in tiles (row stride DP + 4 floats: conflict-free 128-bit loads with lane = candidate); the
// running top-K of a query is one (distance, index) pair per lane, kept sorted with ballot / shuffle insertion.
#include "common.cuh"

namespace gm {
namespace {

constexpr int KNN_WARPS = 8;
constexpr int KNN_THREADS = KNN_WARPS * 32;
#ifndef KNN_QPW64
#define KNN_QPW64 5
#endif
#ifndef KNN_QPW32
#define KNN_QPW32 2
#endif
#ifndef QSMEM_ENABLED
#define QSMEM_ENABLED 1
#endif
#ifndef KNN_TILE_ROWS
#define KNN_TILE_ROWS 128
#endif
constexpr int KNN_TILE = KNN_TILE_ROWS;   // candidates per shared-memory tile (dynamic shared memory: 2 x TILE x (DP + 4) floats)
constexpr int KNN_MAX_K = 32;      // list entries per query = one per lane

__device__ __forceinline__ void cp_async16(float *smem_dst, const float *gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src)
                 : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

__device__ __forceinline__ int event_of(const int *__restrict__ ptr, int B, int i) {
    int lo = 0, hi = B;            // last e with ptr[e] <= i
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(ptr + mid) <= i) lo = mid; else hi = mid;
    }
    return lo;
}

template <int DP, int QPW, int CPL>
__global__ void __launch_bounds__(KNN_THREADS) k_knn_search(const float *__restrict__ x, int stride, int dim,
                                                            const int *__restrict__ ptr, int Ncap, int B, int K, int loop,
                                                            int *__restrict__ nbr, int *__restrict__ deg,
                                                            const int *__restrict__ n_dev) {
    constexpr int RS = DP + 4;                       // shared row stride (floats)
    constexpr int QB = KNN_WARPS * QPW;              // queries per CTA
    constexpr int TILE = DP >= 64 ? (KNN_TILE * 3) / 4 : KNN_TILE;     // 96 rows at 64-d: 4 resident CTAs per SM
    extern __shared__ __align__(16) float s_tile_raw[];      // [2][TILE * RS]: tile t+1 lands (cp.async) while tile t is searched
    float *const s_tile[2] = {s_tile_raw, s_tile_raw + TILE * RS};
    // QS: the warp's QPW queries live in SHARED memory and are read with broadcast LDS.128 (one 16-byte wavefront),
    // the candidate row is read once per 4 dimensions (a 512-byte, 4-wavefront LDS.128) and shared by all QPW
    // queries.  With one query per warp and its coordinates in registers the search was bound by the shared-memory
    // pipe (64 cycles of candidate rows per 32 FP cycles); with 8 queries per warp it is FP-bound.
    constexpr bool QS = QSMEM_ENABLED && QPW >= 4 && DP >= 32;
    float *const s_q = s_tile_raw + 2 * TILE * RS;          // [KNN_WARPS * QPW][DP] (QS only)
    const uint32_t sq_addr = (uint32_t)__cvta_generic_to_shared(s_q + (threadIdx.x >> 5) * QPW * DP);
    pdl_sync();
    // n_dev (optional): the node count lives on the device (pooled DRN graph); rows beyond it get degree 0
    const int N = n_dev ? max(0, min(Ncap, *n_dev)) : Ncap;
    // 16-byte asynchronous copies when rows are whole float4s; otherwise element-wise with zero padding up to DP
    const bool vec = dim == DP && (stride & 3) == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0;
    auto load_tile = [&](int buf, int t0, int rows) {
        float *dst = s_tile[buf];
        if (vec) {
            for (int c = threadIdx.x; c < rows * (DP / 4); c += KNN_THREADS) {
                const int r = c / (DP / 4), d4 = c - r * (DP / 4);
                cp_async16(dst + r * RS + 4 * d4, x + (size_t)(t0 + r) * stride + 4 * d4);
            }
        } else {
            for (int t = threadIdx.x; t < rows * DP; t += KNN_THREADS) {
                const int r = t / DP, d = t - r * DP;
                dst[r * RS + d] = d < dim ? x[(size_t)(t0 + r) * stride + d] : 0.f;
            }
        }
    };
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (long long blk = blockIdx.x; blk * QB < N; blk += gridDim.x) {
        const int q0 = (int)(blk * QB), q1 = min(N, q0 + QB);
        // ---- this warp's queries: coordinates (zero-padded to DP), event, list
        float q[QPW][QS ? 1 : DP];
        int qi[QPW], qe[QPW];
        float ld[QPW], worst[QPW];
        int li[QPW];
#pragma unroll
        for (int u = 0; u < QPW; ++u) {
            qi[u] = q0 + warp * QPW + u;
            const bool live = qi[u] < N;
            qe[u] = live ? event_of(ptr, B, qi[u]) : -1;
            if constexpr (QS) {
                __syncwarp();
                for (int d = lane; d < DP; d += 32)
                    s_q[(warp * QPW + u) * DP + d] = (live && d < dim) ? __ldg(x + (size_t)qi[u] * stride + d) : 0.f;
                __syncwarp();
            } else {
#pragma unroll
                for (int d = 0; d < DP; ++d) q[u][d] = (live && d < dim) ? __ldg(x + (size_t)qi[u] * stride + d) : 0.f;
            }
            ld[u] = INFINITY; li[u] = -1; worst[u] = INFINITY;
        }
        const int e_first = event_of(ptr, B, q0), e_last = event_of(ptr, B, q1 - 1);
        for (int e = e_first; e <= e_last; ++e) {
            const int c0 = __ldg(ptr + e), c1 = __ldg(ptr + e + 1);
            if (c0 >= c1) continue;
            __syncthreads();                         // both buffers free (previous event / query block fully searched)
            load_tile(0, c0, min(TILE, c1 - c0));
            int buf = 0;
            for (int t0 = c0; t0 < c1; t0 += TILE, buf ^= 1) {
                const int rows = min(TILE, c1 - t0);
                cp_async_wait_all();
                __syncthreads();                     // tile `buf` visible to all; everyone is done with tile `buf ^ 1`
                if (t0 + TILE < c1) load_tile(buf ^ 1, t0 + TILE, min(TILE, c1 - t0 - TILE));
                // CPL candidates per lane and step: CPL independent fmaf chains per query hide the FFMA latency at the
                // low occupancy the D = 64 variant runs at (query coordinates fill the register file)
                for (int j0 = 0; j0 < rows; j0 += 32 * CPL) {
                    float acc[CPL][QPW];
                    bool valid[CPL];
                    const float4 *row[CPL];
#pragma unroll
                    for (int g = 0; g < CPL; ++g) {
                        const int j = j0 + 32 * g + lane;
                        valid[g] = j < rows;
                        row[g] = reinterpret_cast<const float4 *>(s_tile[buf] + (valid[g] ? j : 0) * RS);
#pragma unroll
                        for (int u = 0; u < QPW; ++u) acc[g][u] = 0.f;
                    }
#pragma unroll
                    for (int d4 = 0; d4 < DP / 4; ++d4) {
#pragma unroll
                        for (int g = 0; g < CPL; ++g) {
                            const float4 c = row[g][d4];
#pragma unroll
                            for (int u = 0; u < QPW; ++u) {
                                if constexpr (QS) {
                                    float4 qv;                              // volatile: keep the loop-invariant load IN the loop
                                    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                                                 : "=f"(qv.x), "=f"(qv.y), "=f"(qv.z), "=f"(qv.w)
                                                 : "r"(sq_addr + (uint32_t)(u * DP * 4 + 16 * d4)));
                                    const float2 d01 = fsub2(make_float2(c.x, c.y), make_float2(qv.x, qv.y));
                                    const float2 d23 = fsub2(make_float2(c.z, c.w), make_float2(qv.z, qv.w));
                                    acc[g][u] = __fmaf_rn(d01.x, d01.x, acc[g][u]);
                                    acc[g][u] = __fmaf_rn(d01.y, d01.y, acc[g][u]);
                                    acc[g][u] = __fmaf_rn(d23.x, d23.x, acc[g][u]);
                                    acc[g][u] = __fmaf_rn(d23.y, d23.y, acc[g][u]);
                                    continue;
                                }
#ifndef KNN_SCALAR_SUB
                                // the two subtractions of a dimension pair as ONE packed FADD2 (bit-identical halves);
                                // the fmaf chain stays scalar and in dimension order
                                const float2 d01 = fsub2(make_float2(c.x, c.y), make_float2(q[u][4 * d4 + 0], q[u][4 * d4 + 1]));
                                const float2 d23 = fsub2(make_float2(c.z, c.w), make_float2(q[u][4 * d4 + 2], q[u][4 * d4 + 3]));
                                acc[g][u] = __fmaf_rn(d01.x, d01.x, acc[g][u]);
                                acc[g][u] = __fmaf_rn(d01.y, d01.y, acc[g][u]);
                                acc[g][u] = __fmaf_rn(d23.x, d23.x, acc[g][u]);
                                acc[g][u] = __fmaf_rn(d23.y, d23.y, acc[g][u]);
#else
                                float df;
                                df = __fsub_rn(c.x, q[u][4 * d4 + 0]); acc[g][u] = __fmaf_rn(df, df, acc[g][u]);
                                df = __fsub_rn(c.y, q[u][4 * d4 + 1]); acc[g][u] = __fmaf_rn(df, df, acc[g][u]);
                                df = __fsub_rn(c.z, q[u][4 * d4 + 2]); acc[g][u] = __fmaf_rn(df, df, acc[g][u]);
                                df = __fsub_rn(c.w, q[u][4 * d4 + 3]); acc[g][u] = __fmaf_rn(df, df, acc[g][u]);
#endif
                            }
                        }
                    }
#pragma unroll
                    for (int g = 0; g < CPL; ++g) {          // ascending candidate index: group 0 before group 1
#pragma unroll
                        for (int u = 0; u < QPW; ++u) {
                            unsigned pass = __ballot_sync(0xffffffffu, valid[g] && qe[u] == e && acc[g][u] < worst[u]);
                            while (pass) {
                                const int b = __ffs(pass) - 1;
                                pass &= pass - 1;
                                const float cd = __shfl_sync(0xffffffffu, acc[g][u], b);
                                if (!(cd < worst[u])) continue;          // the worst entry moved since the ballot
                                const int ci = t0 + j0 + 32 * g + b;
                                const int pos = __popc(__ballot_sync(0xffffffffu, ld[u] <= cd));   // ties: earlier index first
                                const float upd = __shfl_up_sync(0xffffffffu, ld[u], 1);
                                const int upi = __shfl_up_sync(0xffffffffu, li[u], 1);
                                if (lane > pos) { ld[u] = upd; li[u] = upi; }
                                else if (lane == pos) { ld[u] = cd; li[u] = ci; }
                                // (lanes >= K keep what falls off the end: never <= cd again, never written out)
                                worst[u] = __shfl_sync(0xffffffffu, ld[u], K - 1);
                            }
                        }
                    }
                }
            }
        }
        // ---- write out: self dropped when !loop, survivors packed to the front in (distance, index) order
#pragma unroll
        for (int u = 0; u < QPW; ++u) {
            if (qi[u] >= N) continue;
            const bool keep = lane < K && li[u] >= 0 && (loop || li[u] != qi[u]);
            const unsigned m = __ballot_sync(0xffffffffu, keep);
            const int rank = __popc(m & ((1u << lane) - 1u));
            int *out = nbr + (size_t)qi[u] * K;
            if (keep) out[rank] = li[u];
            const int n = __popc(m);
            if (lane >= n && lane < K) out[lane] = -1;
            if (lane == 0) deg[qi[u]] = n;
        }
    }
    for (int i = N + blockIdx.x * KNN_THREADS + threadIdx.x; i <= Ncap; i += gridDim.x * KNN_THREADS) deg[i] = 0;
}

__global__ void __launch_bounds__(256) k_knn_fill(const int *__restrict__ nbr, const int *__restrict__ rowptr, int N, int K,
                                                  int *__restrict__ col, long long capacity,
                                                  long long *__restrict__ ei64) {
    pdl_sync();
    const long long total = (long long)N * K;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        const int i = (int)(t / K), s = (int)(t - (long long)i * K);
        const int a = rowptr[i], n = rowptr[i + 1] - a;
        if (s >= n) continue;
        const long long p = (long long)a + s;
        if (p >= capacity) continue;
        const int j = nbr[t];
        if (col) col[p] = j;
        if (ei64) { ei64[p] = j; ei64[capacity + p] = i; }      // row 0 = source (neighbour), row 1 = target (query)
    }
}

struct KnnWs {
    int *nbr;       // [N, K]
    int *tiles;     // scan tile totals
};

size_t knn_layout(int N, int K, char *base, KnnWs *w) {
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; };
    const size_t o_nbr = take((size_t)max(N, 1) * K * sizeof(int));
    const size_t o_tiles = take(((size_t)(N + 1) / 1024 + 2) * sizeof(int));
    if (w) { w->nbr = (int *)(base + o_nbr); w->tiles = (int *)(base + o_tiles); }
    return off;
}

}  // namespace

int exclusive_scan_inplace(int *data, int n, int *tile_totals, int *total_out, cudaStream_t st);

namespace {

template <int DP, int QPW, int CPL = 1>
int launch_search(const float *x, int stride, int dim, const int *ptr, int N, int B, int K, int loop, int *nbr, int *deg,
                  const int *n_dev, cudaStream_t st) {
    const int qb = KNN_WARPS * QPW;
    const int grid = persistent_grid(((long long)N + qb - 1) / qb, 16);
    constexpr int TILE = DP >= 64 ? (KNN_TILE * 3) / 4 : KNN_TILE;
    constexpr size_t smem = (size_t)(2 * TILE * (DP + 4) + KNN_WARPS * QPW * DP) * sizeof(float);
    static bool attr = false;
    if (!attr) {
        if (smem > 48 * 1024)
            GM_CUDA_TRY(cudaFuncSetAttribute(k_knn_search<DP, QPW, CPL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr = true;
    }
    GM_LAUNCH_PDL((k_knn_search<DP, QPW, CPL>), grid, KNN_THREADS, smem, st, x, stride, dim, ptr, N, B, K, loop, nbr, deg, n_dev);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

}  // namespace

// the search alone (padded neighbour table nbr [N,K] + deg [N+1], no scan): what the fused DRN forward consumes
int knn_search_table(const float *x, int stride, int dim, const int *ptr, int N, int B, int K, int loop, int *nbr, int *deg,
                     const int *n_dev, cudaStream_t st) {
    if (dim <= 4) return launch_search<4, 4>(x, stride, dim, ptr, N, B, K, loop, nbr, deg, n_dev, st);
    if (dim <= 8) return launch_search<8, 4>(x, stride, dim, ptr, N, B, K, loop, nbr, deg, n_dev, st);
    if (dim <= 16) return launch_search<16, 2>(x, stride, dim, ptr, N, B, K, loop, nbr, deg, n_dev, st);
    if (dim <= 32) return launch_search<32, KNN_QPW32>(x, stride, dim, ptr, N, B, K, loop, nbr, deg, n_dev, st);
    return launch_search<64, KNN_QPW64, 1>(x, stride, dim, ptr, N, B, K, loop, nbr, deg, n_dev, st);
}

}  // namespace gm

using namespace gm;

extern "C" {

size_t gm_knn_graph_ws_bytes(int N, int k, int loop) {
    const int K = k + (loop ? 0 : 1);
    return knn_layout(N < 0 ? 0 : N, K < 1 ? 1 : K, nullptr, nullptr);
}

int gm_knn_graph_count(const float *x, int x_stride, int dim, const int *ptr, int N, int B, int k, int loop, void *ws,
                       size_t ws_bytes, int *rowptr_out, void *stream) {
    if (N < 0 || B < 0 || !rowptr_out || !ws || (N > 0 && (!x || !ptr || B < 1))) return GM_ERR_INVALID_ARGUMENT;
    const int K = k + (loop ? 0 : 1);
    if (k < 1 || K > KNN_MAX_K || dim < 1 || dim > 64 || x_stride < dim) return GM_ERR_UNSUPPORTED;
    if (ws_bytes < knn_layout(N, K, nullptr, nullptr)) return GM_ERR_CAPACITY;
    cudaStream_t st = as_stream(stream);
    if (N == 0) {
        GM_CUDA_TRY(cudaMemsetAsync(rowptr_out, 0, sizeof(int), st));
        return GM_OK;
    }
    KnnWs w;
    knn_layout(N, K, (char *)ws, &w);
    GM_TRY(knn_search_table(x, x_stride, dim, ptr, N, B, K, loop, w.nbr, rowptr_out, nullptr, st));
    GM_TRY(exclusive_scan_inplace(rowptr_out, N + 1, w.tiles, nullptr, st));
    return GM_OK;
}

int gm_knn_graph_fill(int N, int k, int loop, const void *ws, size_t ws_bytes, const int *rowptr, int *col_out,
                      long long col_capacity, int64_t *edge_index64_out, void *stream) {
    if (N < 0 || !ws || !rowptr || col_capacity < 0) return GM_ERR_INVALID_ARGUMENT;
    const int K = k + (loop ? 0 : 1);
    if (k < 1 || K > KNN_MAX_K) return GM_ERR_UNSUPPORTED;
    if (ws_bytes < knn_layout(N, K, nullptr, nullptr)) return GM_ERR_CAPACITY;
    if (N == 0 || col_capacity == 0) return GM_OK;
    KnnWs w;
    knn_layout(N, K, (char *)ws, &w);
    cudaStream_t st = as_stream(stream);
    const int grid = persistent_grid(((long long)N * K + 255) / 256, 8);
    GM_LAUNCH_PDL(k_knn_fill, grid, 256, 0, st, (const int *)w.nbr, rowptr, N, K, col_out, col_capacity,
                  (long long *)edge_index64_out);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

}  // extern "C"
