// a2. Fixed-radius neighbour search in (eta, phi): cell-binned per-event search with warp-level
// compaction into a CSR sorted by (target, source).  Replaces torch_cluster.radius_graph as
// called at train.py:48 / evaluate.py:88 of the reference (brute force, one thread per query,
// O(sum n_e^2), N*255 int64 temporaries).  Bit-exact to that edge set (see oracle/radius_bruteforce.c).
//
// Pipeline (all on `stream`, no host sync):
//   bbox -> cell histogram -> per-event cell scan -> cell scatter (sorted float4 copy of the points)
//   -> degree count (warp per target, 3x3 cell neighbourhood) -> exclusive scan -> fill.
// The fill kernel marks hits in a per-warp shared-memory bitmask indexed by the event-local source
// index and then emits the set bits in order: the row comes out sorted by source index and the
// max_num_neighbors cap keeps the LOWEST indices, exactly like the reference's sequential scan,
// without any sort and independent of the (atomic, hence arbitrary) order inside a cell.
#include "geo_layout.cuh"
#include <cstdio>
#include <mutex>

namespace gm {
namespace {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

struct Workspace {
    int *bbox;        // [4] ordered-int min0,max0,min1,max1
    int *cell_cnt;    // [B*CMAX]
    int *cell_start;  // [B*(CMAX+1)]
    float4 *sorted;   // [N] (eta, phi, local index bits, cell id bits) grouped by event, cell
    int *scan_tot;    // [ceil((N+1)/SCAN_TILE)]
    int *stats;       // [4] E, max degree, #capped, flags
};

size_t ws_layout(int N, int B, char *base, Workspace *w) {
    size_t off = 0;
    auto take = [&](size_t bytes) {
        size_t o = off;
        off = align_up(off + bytes, 256);
        return base ? base + o : nullptr;
    };
    int *bbox = (int *)take(4 * sizeof(int));
    int *cnt = (int *)take((size_t)B * CMAX * sizeof(int));
    int *start = (int *)take((size_t)B * (CMAX + 1) * sizeof(int));
    float4 *sorted = (float4 *)take((size_t)(N > 0 ? N : 1) * sizeof(float4));
    int *tot = (int *)take(((size_t)(N + 1) / SCAN_TILE + 2) * sizeof(int));
    int *stats = (int *)take(4 * sizeof(int));
    if (w) *w = Workspace{bbox, cnt, start, sorted, tot, stats};
    return off;
}

// ------------------------------------------------------------------------------------------
// bbox / stats reset and the cell histogram cleared in one launch (a memset node here would also break the
// programmatic-dependent-launch chain of the kernels that follow)
__global__ void __launch_bounds__(256) k_init(int *bbox, int *stats, int4 *__restrict__ cell_cnt4, int n4, GeoInit gi,
                                              const int *__restrict__ ptr) {
    pdl_sync();
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        bbox[0] = 0x7fffffff; bbox[1] = (int)0x80000000;
        bbox[2] = 0x7fffffff; bbox[3] = (int)0x80000000;
        stats[0] = 0; stats[1] = 0; stats[2] = 0; stats[3] = 0;
    }
    if (gi.meta) geo_init_body(gi, ptr, stats);          // gm_geo_build's per-build reset rides along (thread 0 of CTA 0 again)
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += gridDim.x * blockDim.x) cell_cnt4[i] = make_int4(0, 0, 0, 0);
}

__global__ void __launch_bounds__(256) k_bbox(const float *__restrict__ pos, int stride, int dim, int N, int *bbox) {
    pdl_sync();
    int lo0 = 0x7fffffff, hi0 = (int)0x80000000, lo1 = lo0, hi1 = hi0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
        float a = pos[(size_t)i * stride];
        if (isfinite(a)) { int o = float_to_ordered(a); lo0 = min(lo0, o); hi0 = max(hi0, o); }
        if (dim >= 2) {
            float b = pos[(size_t)i * stride + 1];
            if (isfinite(b)) { int o = float_to_ordered(b); lo1 = min(lo1, o); hi1 = max(hi1, o); }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo0 = min(lo0, __shfl_xor_sync(0xffffffffu, lo0, o));
        hi0 = max(hi0, __shfl_xor_sync(0xffffffffu, hi0, o));
        lo1 = min(lo1, __shfl_xor_sync(0xffffffffu, lo1, o));
        hi1 = max(hi1, __shfl_xor_sync(0xffffffffu, hi1, o));
    }
    if ((threadIdx.x & 31) == 0) {  // integer min/max: order-independent, deterministic
        atomicMin(bbox + 0, lo0); atomicMax(bbox + 1, hi0);
        atomicMin(bbox + 2, lo1); atomicMax(bbox + 3, hi1);
    }
}

__global__ void __launch_bounds__(256) k_cell_count(const float *__restrict__ pos, int stride, int dim,
                                                    const int *__restrict__ ptr, int N, int B, float r,
                                                    int phi_wrap, const int *__restrict__ bbox,
                                                    int *__restrict__ cell_cnt) {
    pdl_sync();
    CellGrid g = make_grid(bbox, r, dim, phi_wrap);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
        int e = find_event(ptr, B, i);
        float a = pos[(size_t)i * stride];
        float b = dim >= 2 ? pos[(size_t)i * stride + 1] : 0.f;
        int c = cell_coord(a, g.min0, g.inv0, g.g0) * g.g1 + cell_coord(b, g.min1, g.inv1, g.g1);
        atomicAdd(cell_cnt + (size_t)e * CMAX + c, 1);
    }
}

// one CTA per event: cell_start[e][c] = ptr[e] + exclusive_scan(cell_cnt[e][:]) ; CMAX = 1024 = blockDim
__global__ void __launch_bounds__(CMAX) k_cell_scan(const int *__restrict__ ptr, const int *__restrict__ cell_cnt,
                                                   int *__restrict__ cell_start) {
    pdl_sync();
    __shared__ int warp_tot[32];
    const int e = blockIdx.x, t = threadIdx.x, lane = t & 31, w = t >> 5;
    int v = cell_cnt[(size_t)e * CMAX + t];
    int tot;
    int ex = warp_excl_scan_int(v, lane, tot);
    if (lane == 31) warp_tot[w] = tot;
    __syncthreads();
    if (w == 0) {
        int wt = warp_tot[lane], dummy;
        int wex = warp_excl_scan_int(wt, lane, dummy);
        warp_tot[lane] = wex;
    }
    __syncthreads();
    int base = ptr[e];
    int s = base + warp_tot[w] + ex;
    cell_start[(size_t)e * (CMAX + 1) + t] = s;
    if (t == CMAX - 1) cell_start[(size_t)e * (CMAX + 1) + CMAX] = s + v;
}

__global__ void __launch_bounds__(256) k_cell_fill(const float *__restrict__ pos, int stride, int dim,
                                                   const int *__restrict__ ptr, int N, int B, float r,
                                                   int phi_wrap, const int *__restrict__ bbox,
                                                   int *__restrict__ cell_cnt, const int *__restrict__ cell_start,
                                                   float4 *__restrict__ sorted) {
    pdl_sync();
    CellGrid g = make_grid(bbox, r, dim, phi_wrap);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
        int e = find_event(ptr, B, i);
        float a = pos[(size_t)i * stride];
        float b = dim >= 2 ? pos[(size_t)i * stride + 1] : 0.f;
        int c = cell_coord(a, g.min0, g.inv0, g.g0) * g.g1 + cell_coord(b, g.min1, g.inv1, g.g1);
        // counts are consumed back to zero: the histogram buffer is clean for the next call
        int k = atomicSub(cell_cnt + (size_t)e * CMAX + c, 1) - 1;
        int slot = cell_start[(size_t)e * (CMAX + 1) + c] + k;
        sorted[slot] = make_float4(a, b, __int_as_float(i - ptr[e]), __int_as_float(c));
    }
}

// Walk the (<= 6) contiguous runs of the sorted array that cover the 3x3 cell neighbourhood.
template <typename F>
__device__ __forceinline__ void for_each_candidate(const CellGrid &g, const int *__restrict__ cs, int cell,
                                                   int phi_wrap, int lane, F &&f) {
    const int ce = cell / g.g1, cp = cell - ce * g.g1;
    for (int de = -1; de <= 1; ++de) {
        int ce2 = ce + de;
        if (ce2 < 0 || ce2 >= g.g0) continue;
        const int *row = cs + ce2 * g.g1;
        int lo = cp - 1, hi = cp + 1;
        int runs_lo[2], runs_hi[2], nruns = 1;
        if (phi_wrap && g.g1 >= 3) {
            if (lo < 0) { runs_lo[1] = row[g.g1 - 1]; runs_hi[1] = row[g.g1]; nruns = 2; lo = 0; }
            else if (hi >= g.g1) { runs_lo[1] = row[0]; runs_hi[1] = row[1]; nruns = 2; hi = g.g1 - 1; }
        } else if (phi_wrap) {  // fewer than 3 phi cells: every phi cell is a neighbour
            lo = 0; hi = g.g1 - 1;
        }
        lo = lo < 0 ? 0 : lo;
        hi = hi >= g.g1 ? g.g1 - 1 : hi;
        runs_lo[0] = row[lo]; runs_hi[0] = row[hi + 1];
        for (int rr = 0; rr < nruns; ++rr)
            for (int idx = runs_lo[rr] + lane; idx < runs_hi[rr]; idx += 32) f(idx);
    }
}

__global__ void __launch_bounds__(256) k_degree(const int *__restrict__ ptr, int N, int B, float r, float r2,
                                                int dim, int cap, int loop, int phi_wrap,
                                                const int *__restrict__ bbox, const int *__restrict__ cell_start,
                                                const float4 *__restrict__ sorted, int *__restrict__ deg,
                                                int *__restrict__ stats) {
    const CellGrid g = make_grid(bbox, r, dim, phi_wrap);
    const int lane = threadIdx.x & 31;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    int local_max = 0, local_capped = 0;
    for (int s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; s < N; s += warps) {
        const int e = find_event(ptr, B, s);
        const int base = __ldg(ptr + e);
        const float4 t = sorted[s];
        const int tl = __float_as_int(t.z), cell = __float_as_int(t.w);
        int total = 0, lower = 0, self_hit = 0;
        for_each_candidate(g, cell_start + (size_t)e * (CMAX + 1), cell, phi_wrap, lane, [&](int idx) {
            float4 c = sorted[idx];
            int cl = __float_as_int(c.z);
            bool hit = dist2(t.x, t.y, c.x, c.y, dim, phi_wrap) < r2;
            total += hit;
            lower += hit && (cl < tl);
            self_hit |= hit && (cl == tl);
        });
        total = warp_sum_int(total);
        lower = warp_sum_int(lower);
        self_hit = __any_sync(0xffffffffu, self_hit);
        int d;
        if (loop) d = min(total, cap);
        else d = min(total, cap + 1) - ((self_hit && lower < cap + 1) ? 1 : 0);
        if (lane == 0) {
            deg[base + tl] = d;
            local_max = max(local_max, total);
            local_capped += total > (loop ? cap : cap + 1);
        }
    }
    if (lane == 0) {
        if (local_max) atomicMax(stats + 1, local_max);
        if (local_capped) { atomicAdd(stats + 2, local_capped); atomicOr(stats + 3, GM_FLAG_DEGREE_CAPPED); }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) deg[N] = 0;
}

// exclusive scan, phase 1: per-tile totals
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_totals(const int *__restrict__ in, int n, int *__restrict__ tot) {
    __shared__ int sm[SCAN_THREADS / 32];
    int base = blockIdx.x * SCAN_TILE;
    int v = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        int i = base + k * SCAN_THREADS + threadIdx.x;
        if (i < n) v += in[i];
    }
    v = warp_sum_int(v);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        int s = 0;
        for (int k = 0; k < SCAN_THREADS / 32; ++k) s += sm[k];
        tot[blockIdx.x] = s;
    }
}

// phase 2: every tile sums the totals of the tiles before it (no inter-CTA waiting), then scans in place
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_apply(int *__restrict__ data, int n, const int *__restrict__ tot,
                                                            int *__restrict__ total_out) {
    __shared__ int sm[SCAN_THREADS / 32];
    __shared__ int s_off;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int acc = 0;
    for (int b = threadIdx.x; b < (int)blockIdx.x; b += SCAN_THREADS) acc += tot[b];
    acc = warp_sum_int(acc);
    if (lane == 0) sm[w] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        int s = 0;
        for (int k = 0; k < SCAN_THREADS / 32; ++k) s += sm[k];
        s_off = s;
    }
    __syncthreads();
    const int off = s_off;
    __syncthreads();
    // each thread owns SCAN_ITEMS consecutive elements
    int base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    int v[SCAN_ITEMS], sum = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) { v[k] = (base + k < n) ? data[base + k] : 0; sum += v[k]; }
    int wtot;
    int ex = warp_excl_scan_int(sum, lane, wtot);
    if (lane == 31) sm[w] = wtot;
    __syncthreads();
    int woff = 0;
    for (int k = 0; k < w; ++k) woff += sm[k];
    int run = off + woff + ex;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        if (base + k < n) data[base + k] = run;
        run += v[k];
    }
    if (total_out && base <= n - 1 && n - 1 < base + SCAN_ITEMS) {
        // the thread owning the last element (deg[N] == 0 => exclusive value there is the grand total)
        int last = off + woff + ex;
        for (int k = 0; k < n - 1 - base; ++k) last += v[k];
        *total_out = last;
    }
}

__global__ void __launch_bounds__(256) k_fill(const int *__restrict__ ptr, int N, int B, float r, float r2, int dim,
                                              int cap, int loop, int phi_wrap, int mask_words,
                                              const int *__restrict__ bbox, const int *__restrict__ cell_start,
                                              const float4 *__restrict__ sorted, const int *__restrict__ rowptr,
                                              int *__restrict__ col, long long capacity,
                                              long long *__restrict__ ei64, int *__restrict__ stats) {
    extern __shared__ unsigned int smask[];
    const CellGrid g = make_grid(bbox, r, dim, phi_wrap);
    const int lane = threadIdx.x & 31;
    unsigned int *mask = smask + (size_t)(threadIdx.x >> 5) * mask_words;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    const int cap_eff = loop ? cap : cap + 1;
    for (int s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; s < N; s += warps) {
        const int e = find_event(ptr, B, s);
        const int base = __ldg(ptr + e);
        const int n_e = __ldg(ptr + e + 1) - base;
        const int words = (n_e + 31) >> 5;
        if (words > mask_words) {  // host promised a smaller max_event_nodes
            if (lane == 0) atomicOr(stats + 3, GM_FLAG_EVENT_TOO_LARGE);
            continue;
        }
        const float4 t = sorted[s];
        const int tl = __float_as_int(t.z), cell = __float_as_int(t.w);
        for (int w = lane; w < words; w += 32) mask[w] = 0u;
        __syncwarp();
        int lower = 0, self_hit = 0;
        for_each_candidate(g, cell_start + (size_t)e * (CMAX + 1), cell, phi_wrap, lane, [&](int idx) {
            float4 c = sorted[idx];
            int cl = __float_as_int(c.z);
            bool hit = dist2(t.x, t.y, c.x, c.y, dim, phi_wrap) < r2;
            if (hit) atomicOr(mask + (cl >> 5), 1u << (cl & 31));
            lower += hit && (cl < tl);
            self_hit |= hit && (cl == tl);
        });
        __syncwarp();
        lower = warp_sum_int(lower);
        self_hit = __any_sync(0xffffffffu, self_hit);
        const int tg = base + tl;
        const int out = rowptr[tg];
        const int deg = rowptr[tg + 1] - out;
        if ((long long)out + deg > capacity) {
            if (lane == 0) atomicOr(stats + 3, GM_FLAG_EDGE_OVERFLOW);
            continue;
        }
        int rank_base = 0;
        for (int w0 = 0; w0 < words && rank_base < cap_eff; w0 += 32) {
            unsigned int word = (w0 + lane < words) ? mask[w0 + lane] : 0u;
            int tot;
            int rk = rank_base + warp_excl_scan_int(__popc(word), lane, tot);
            while (word) {
                int b = __ffs(word) - 1;
                word &= word - 1;
                int l = ((w0 + lane) << 5) + b;
                if (rk < cap_eff && (loop || l != tl)) {
                    int pos = loop ? rk : rk - ((self_hit && rk > lower) ? 1 : 0);
                    col[out + pos] = base + l;
                    if (ei64) { ei64[out + pos] = base + l; ei64[capacity + out + pos] = tg; }
                }
                ++rk;
            }
            rank_base += tot;
        }
        __syncwarp();
    }
}

__global__ void __launch_bounds__(256) k_prepare_etaphi(const float *__restrict__ x, int stride, int N,
                                                        float2 *__restrict__ out) {
    pdl_sync();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
        const float *row = x + (size_t)i * stride;
        out[i] = make_float2(row[3], atan2f(row[1], row[0]));
    }
}

__global__ void __launch_bounds__(256) k_batch_to_ptr(const long long *__restrict__ batch, int N, int B,
                                                      int *__restrict__ ptr) {
    // ptr[b] = first node index whose event id is >= b  (empty events allowed anywhere)
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i <= N; i += gridDim.x * blockDim.x) {
        long long prev = i == 0 ? -1 : batch[i - 1];
        long long cur = i == N ? (long long)B : batch[i];
        if (cur > B) cur = B;
        for (long long b = prev + 1; b <= cur; ++b) ptr[b] = i;
    }
}

int run_common(const float *pos, int pos_stride, int dim, const int *ptr, int N, int B, double r,
               int phi_wrap, const Workspace &w, cudaStream_t st, const GeoInit *geo_init = nullptr,
               bool prologue_only = false) {
    const float rf = (float)r;
    const int n4 = B * (CMAX / 4);                            // cell_cnt is 256-byte aligned in both workspaces, CMAX % 4 == 0
    GeoInit gi{};
    if (geo_init) gi = *geo_init;
    const int init_work = (n4 > N ? n4 : N) + 255;
    GM_LAUNCH_PDL(k_init, persistent_grid(init_work / 256 > 0 ? init_work / 256 : 1, 8), 256, 0, st, w.bbox, w.stats,
                  reinterpret_cast<int4 *>(w.cell_cnt), n4, gi, ptr);
    GM_LAUNCH_CHECK();
    if (N == 0) return GM_OK;
    const int g = persistent_grid((N + 255) / 256, 8);
    GM_LAUNCH_PDL(k_bbox, g, 256, 0, st, pos, pos_stride, dim, N, w.bbox);
    GM_LAUNCH_CHECK();
    if (prologue_only) return GM_OK;                          // the caller bins the cells itself (geo: one CTA per event)
    GM_LAUNCH_PDL(k_cell_count, g, 256, 0, st, pos, pos_stride, dim, ptr, N, B, rf, phi_wrap, w.bbox, w.cell_cnt);
    GM_LAUNCH_CHECK();
    GM_LAUNCH_PDL(k_cell_scan, B, CMAX, 0, st, ptr, w.cell_cnt, w.cell_start);
    GM_LAUNCH_CHECK();
    GM_LAUNCH_PDL(k_cell_fill, g, 256, 0, st, pos, pos_stride, dim, ptr, N, B, rf, phi_wrap, w.bbox, w.cell_cnt, w.cell_start,
                  w.sorted);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

int check_args(const float *pos, int dim, const int *ptr, int N, int B, double r, int max_nn, void *ws,
               size_t ws_bytes) {
    if (N < 0 || B < 0 || (N > 0 && (!pos || !ptr || B == 0)) || !ws) return GM_ERR_INVALID_ARGUMENT;
    if (!(r > 0.0) || max_nn < 1) return GM_ERR_INVALID_ARGUMENT;
    if (dim < 1 || dim > 2) return GM_ERR_UNSUPPORTED;
    if (ws_bytes < ws_layout(N, B, nullptr, nullptr)) return GM_ERR_CAPACITY;
    return GM_OK;
}

}  // namespace

int bin_points(const float *pos, int pos_stride, int dim, const int *ptr, int N, int B, float r, int phi_wrap,
               int *bbox, int *cell_cnt, int *cell_start, float4 *sorted, int *stats, cudaStream_t st,
               const GeoInit *geo_init, bool prologue_only) {
    Workspace w{bbox, cell_cnt, cell_start, sorted, nullptr, stats};
    return run_common(pos, pos_stride, dim, ptr, N, B, (double)r, phi_wrap, w, st, geo_init, prologue_only);
}

// ------------------------------------------------------------------------------------------
// Scratch of gm_radius_graph_static: lives in the library's image on every GPU that loads it (until exit), so the
// entry point never allocates.  Shared by ALL calls: two calls that overlap on the GPU race on it.
constexpr int STATIC_MAX_NODES = GM_STATIC_MAX_NODES, STATIC_MAX_EVENTS = GM_STATIC_MAX_EVENTS;
constexpr size_t STATIC_SCRATCH_BYTES = 1600u * 1024u;       // >= ws_layout(STATIC_MAX_NODES, STATIC_MAX_EVENTS) = 1,573,888
static __device__ __align__(256) unsigned char g_static_scratch[STATIC_SCRATCH_BYTES];

constexpr int STATIC_MAX_DEVICES = 64;
static std::mutex g_static_mu;
static void *g_static_ptr[STATIC_MAX_DEVICES];               // device address of g_static_scratch, per device

// In host code `g_static_scratch` is only a handle: the device address comes from cudaGetSymbolAddress, looked up once
// per device with that device current.
static int static_scratch(void **out) {
    int dev = -1;
    GM_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= STATIC_MAX_DEVICES) return GM_ERR_UNSUPPORTED;
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_ptr[dev]) {
        void *p = nullptr;
        GM_CUDA_TRY(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p) return GM_ERR_CUDA;
        g_static_ptr[dev] = p;
    }
    *out = g_static_ptr[dev];
    return GM_OK;
}

int exclusive_scan_inplace(int *data, int n, int *tile_totals, int *total_out, cudaStream_t st) {
    if (n <= 0) return GM_OK;
    int tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    k_scan_totals<<<tiles, SCAN_THREADS, 0, st>>>(data, n, tile_totals);
    GM_LAUNCH_CHECK();
    k_scan_apply<<<tiles, SCAN_THREADS, 0, st>>>(data, n, tile_totals, total_out);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

}  // namespace gm

using namespace gm;

extern "C" {

size_t gm_radius_graph_ws_bytes(int N, int B) { return ws_layout(N < 0 ? 0 : N, B < 0 ? 0 : B, nullptr, nullptr); }

int gm_radius_graph_count(const float *pos, int pos_stride, int dim, const int *ptr, int N, int B, double r,
                          int max_num_neighbors, int loop, int phi_wrap, int max_event_nodes, void *ws,
                          size_t ws_bytes, int *rowptr_out, int *stats_out, void *stream) {
    (void)max_event_nodes;
    GM_TRY(check_args(pos, dim, ptr, N, B, r, max_num_neighbors, ws, ws_bytes));
    if (!rowptr_out) return GM_ERR_INVALID_ARGUMENT;
    cudaStream_t st = as_stream(stream);
    Workspace w;
    ws_layout(N, B, (char *)ws, &w);
    TimedRegion tr(GM_TIMER_RADIUS_GRAPH, st);
    GM_TRY(run_common(pos, pos_stride, dim, ptr, N, B, r, phi_wrap, w, st));
    const float rf = (float)r, r2 = (float)(r * r);  // r*r in double, then fp32
    if (N > 0) {
        const int g = persistent_grid(((long long)N * 32 + 255) / 256, 8);
        k_degree<<<g, 256, 0, st>>>(ptr, N, B, rf, r2, dim, max_num_neighbors, loop, phi_wrap, w.bbox,
                                    w.cell_start, w.sorted, rowptr_out, w.stats);
        GM_LAUNCH_CHECK();
        GM_TRY(exclusive_scan_inplace(rowptr_out, N + 1, w.scan_tot, w.stats + 0, st));
    } else {
        GM_CUDA_TRY(cudaMemsetAsync(rowptr_out, 0, sizeof(int), st));
    }
    if (stats_out) GM_CUDA_TRY(cudaMemcpyAsync(stats_out, w.stats, 4 * sizeof(int), cudaMemcpyDeviceToDevice, st));
    return GM_OK;
}

int gm_radius_graph_fill(const float *pos, int pos_stride, int dim, const int *ptr, int N, int B, double r,
                         int max_num_neighbors, int loop, int phi_wrap, int max_event_nodes, void *ws,
                         size_t ws_bytes, const int *rowptr, int *col_out, long long col_capacity,
                         int64_t *edge_index64_out, int *stats_out, void *stream) {
    GM_TRY(check_args(pos, dim, ptr, N, B, r, max_num_neighbors, ws, ws_bytes));
    if (!rowptr || (!col_out && col_capacity > 0) || max_event_nodes < 0) return GM_ERR_INVALID_ARGUMENT;
    if (N == 0) return GM_OK;
    cudaStream_t st = as_stream(stream);
    Workspace w;
    ws_layout(N, B, (char *)ws, &w);  // sorted points / cell starts were left there by _count
    TimedRegion tr(GM_TIMER_RADIUS_GRAPH, st);
    const int mask_words = (max_event_nodes + 31) / 32 + 1;
    int threads = 256;
    while (threads > 32 && (size_t)(threads / 32) * mask_words * 4 > 200 * 1024) threads >>= 1;
    const size_t smem = (size_t)(threads / 32) * mask_words * 4;
    if (smem > 200 * 1024) return GM_ERR_UNSUPPORTED;
    if (smem > 48 * 1024)
        GM_CUDA_TRY(cudaFuncSetAttribute(k_fill, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const float rf = (float)r, r2 = (float)(r * r);
    const int g = persistent_grid(((long long)N * 32 + threads - 1) / threads, 2048 / threads);
    k_fill<<<g, threads, smem, st>>>(ptr, N, B, rf, r2, dim, max_num_neighbors, loop, phi_wrap, mask_words, w.bbox,
                                     w.cell_start, w.sorted, rowptr, col_out, col_capacity,
                                     (long long *)edge_index64_out, w.stats);
    GM_LAUNCH_CHECK();
    if (stats_out) GM_CUDA_TRY(cudaMemcpyAsync(stats_out, w.stats, 4 * sizeof(int), cudaMemcpyDeviceToDevice, st));
    return GM_OK;
}

// The two-call protocol above in ONE call that allocates nothing: the workspace is g_static_scratch.  Same kernels, same
// launch shapes, same results as gm_radius_graph_count + gm_radius_graph_fill.  k_init clears what must start at zero
// (bbox, counters, cell histogram), so the previous call's contents do not matter.
int gm_radius_graph_static(const float *pos, int pos_stride, int dim, const int *ptr, int N, int B, double r,
                           int max_num_neighbors, int loop, int phi_wrap, int max_event_nodes, int *rowptr_out,
                           int *col_out, long long col_capacity, int *stats_out, void *stream) {
    if (N > STATIC_MAX_NODES || B > STATIC_MAX_EVENTS) {
        fprintf(stderr, "gm_radius_graph_static: N=%d / B=%d beyond the static scratch (GM_STATIC_MAX_NODES=%d, "
                        "GM_STATIC_MAX_EVENTS=%d); use gm_radius_graph_count/_fill with a caller workspace\n",
                N, B, STATIC_MAX_NODES, STATIC_MAX_EVENTS);
        return GM_ERR_CAPACITY;
    }
    void *ws = nullptr;
    if (N >= 0 && B >= 0 && ws_layout(N, B, nullptr, nullptr) > STATIC_SCRATCH_BYTES) return GM_ERR_CAPACITY;
    GM_TRY(static_scratch(&ws));
    GM_TRY(check_args(pos, dim, ptr, N, B, r, max_num_neighbors, ws, STATIC_SCRATCH_BYTES));
    if (!rowptr_out || (!col_out && col_capacity > 0) || col_capacity < 0 || max_event_nodes < 0)
        return GM_ERR_INVALID_ARGUMENT;
    cudaStream_t st = as_stream(stream);
    Workspace w;
    ws_layout(N, B, (char *)ws, &w);
    GM_TRY(run_common(pos, pos_stride, dim, ptr, N, B, r, phi_wrap, w, st));
    if (N == 0) {
        GM_CUDA_TRY(cudaMemsetAsync(rowptr_out, 0, sizeof(int), st));
        if (stats_out) GM_CUDA_TRY(cudaMemcpyAsync(stats_out, w.stats, 4 * sizeof(int), cudaMemcpyDeviceToDevice, st));
        return GM_OK;
    }
    const float rf = (float)r, r2 = (float)(r * r);
    const int gd = persistent_grid(((long long)N * 32 + 255) / 256, 8);
    k_degree<<<gd, 256, 0, st>>>(ptr, N, B, rf, r2, dim, max_num_neighbors, loop, phi_wrap, w.bbox, w.cell_start,
                                 w.sorted, rowptr_out, w.stats);
    GM_LAUNCH_CHECK();
    GM_TRY(exclusive_scan_inplace(rowptr_out, N + 1, w.scan_tot, w.stats + 0, st));
    const int mask_words = (max_event_nodes + 31) / 32 + 1;
    int threads = 256;
    while (threads > 32 && (size_t)(threads / 32) * mask_words * 4 > 200 * 1024) threads >>= 1;
    const size_t smem = (size_t)(threads / 32) * mask_words * 4;
    if (smem > 200 * 1024) return GM_ERR_UNSUPPORTED;
    if (smem > 48 * 1024)
        GM_CUDA_TRY(cudaFuncSetAttribute(k_fill, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int gf = persistent_grid(((long long)N * 32 + threads - 1) / threads, 2048 / threads);
    k_fill<<<gf, threads, smem, st>>>(ptr, N, B, rf, r2, dim, max_num_neighbors, loop, phi_wrap, mask_words, w.bbox,
                                      w.cell_start, w.sorted, rowptr_out, col_out, col_capacity, (long long *)nullptr,
                                      w.stats);
    GM_LAUNCH_CHECK();
    if (stats_out) GM_CUDA_TRY(cudaMemcpyAsync(stats_out, w.stats, 4 * sizeof(int), cudaMemcpyDeviceToDevice, st));
    return GM_OK;
}

int gm_prepare_etaphi(const float *x, int x_stride, int N, float *etaphi_out, void *stream) {
    if (N < 0 || (N > 0 && (!x || !etaphi_out)) || x_stride < 4) return GM_ERR_INVALID_ARGUMENT;
    if (N == 0) return GM_OK;
    GM_LAUNCH_PDL(k_prepare_etaphi, persistent_grid((N + 255) / 256, 8), 256, 0, as_stream(stream), x, x_stride, N,
                  (float2 *)etaphi_out);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

int gm_batch_to_ptr(const int64_t *batch, int N, int B, int *ptr_out, void *stream) {
    if (N < 0 || B < 0 || !ptr_out || (N > 0 && !batch)) return GM_ERR_INVALID_ARGUMENT;
    k_batch_to_ptr<<<persistent_grid((N + 256) / 256, 8), 256, 0, as_stream(stream)>>>((const long long *)batch, N, B,
                                                                                       ptr_out);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

}  // extern "C"
