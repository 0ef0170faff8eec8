// Cell-staged EdgeConv with an IMPLICIT radius graph (the fused path of GraphMETTrainer).
//
// The edge list of radius_graph(etaphi, r, batch, loop, max_num_neighbors) (train.py:48) is never written: nodes run
// in INTERNAL order (event, cell, original index), every neighbour of a node lies in the <= 3 contiguous runs that
// cover the 3x3 cells around it, and the kernels below stage those runs ONCE per block of 32 consecutive nodes into
// shared memory with TMA bulk copies (cp.async.bulk + mbarrier) and re-evaluate the edge predicate of geo_layout.cuh
// there.  Against the CSR kernels (k_edge_aggr / k_scatter_s) this removes the col[] traffic (4E bytes written, read
// 3x), fetches each 128-byte row from L2 once per block instead of once per edge, and needs no transposed CSR for
// the backward, capped graph or not.
//
//   gm_geo_build        : bin -> deterministic order inside cells -> node_perm -> cap thresholds (dense cells only)
//   k_geo_aggr          : out_i = P_i + max_{j in N(i)} Y_j  (+ arg-max, + BatchNorm moments)              [a5 fwd]
//   k_geo_scatter       : S_j = sum_{i : arg(i,c) == j} d conv_i[c], BatchNorm backward on the fly          [a5 bwd]
// All reductions run in a fixed order (internal node order): bit-deterministic, no float atomics.
#include "geo_stage.cuh"

namespace gm {
namespace {

#ifndef GEO_SC_THREADS
#define GEO_SC_THREADS 128
#endif
// 128-row stages (115 KB per CTA: ONE CTA of 5 warps per SM).  Timed alone (own CUDA-event pair) the kernel is slower
// than with 32-row stages and 5 CTAs per SM (c2 31 vs 27 us) -- but inside the dependent-launch chain of a step the lean
// footprint leaves room for the next kernel's CTAs to be resident and through their prologue: the STEP is 3 % faster
// (c2 0.456 -> 0.442 ms, c5 1.104 -> 1.076 ms; 32 / 48 / 64 / 96 / 192 / 256 rows and 3-4 stages measured).
#ifndef GEO_SC_CH
#define GEO_SC_CH 128
#endif
constexpr int GEO_THREADS = GEO_SC_THREADS;     // backward kernel: 4 warps
constexpr int GEO_WARPS = GEO_THREADS / 32;
constexpr int AG_CONSUMERS = 8;      // forward kernel: 8 consumer warps (AG_TPW targets of the item each) ...
constexpr int AG_THREADS = (AG_CONSUMERS + 1) * 32;   // ... + 1 producer warp (segments, descriptors, TMA)
constexpr int AG_TPW = GEO_TPI / AG_CONSUMERS;
// Two tunings of the forward aggregation kernel (geo_aggr.inc), picked per launch by the mean event size:
//   ag_big   : 224-row stages x 2, compiled for 3 CTAs per SM (27 warps) -- 5k-candidate events: dense hit lists want warps
//   ag_small : 320-row stages x 2, compiled for 2 CTAs per SM (112 registers) -- DY/TT-sized events: a segment's whole
//              neighbourhood fits one chunk and the arg-max body no longer sits at its register cap
// Measured on whole steps (timers off): c2 0.442 -> 0.435 ms with ag_small, c5 1.076 -> 1.088 ms (hence the switch).
#ifndef GEO_AG_CH
#define GEO_AG_CH 224
#endif
#ifndef GEO_AG_STAGES
#define GEO_AG_STAGES 2
#endif
#ifndef GEO_AG_MINB
#define GEO_AG_MINB 3
#endif
#ifndef GEO_AG_SMALL_EVENT
#define GEO_AG_SMALL_EVENT 3000   // mean nodes per event below which ag_small runs (training kernel)
#endif


// ------------------------------------------------------------------------------------------ build
// Binning of ONE event per CTA (events of up to GEO_BIN_EVENT_MAX nodes): cell histogram in shared memory, scan,
// counting-sort scatter, re-ranking inside the cells by original index (the internal order must be reproducible:
// see k_cell_rank) and the cap pre-filter -- what k_cell_count / k_cell_scan / k_cell_fill / k_cell_rank do in four
// launches with global atomics.  At DY/TT sizes the graph build is a chain of latency-bound launches; this is one.
constexpr int BE_THREADS = 1024;
constexpr int GEO_BIN_EVENT_MAX = 16384, GEO_BIN_EVENT_MIN_B = 128;
__device__ __forceinline__ int cell_of(const CellGrid &g, const float *__restrict__ pos, int stride, int dim, size_t i,
                                       float &a, float &b) {
    a = pos[i * stride];
    b = dim >= 2 ? pos[i * stride + 1] : 0.f;
    return cell_coord(a, g.min0, g.inv0, g.g0) * g.g1 + cell_coord(b, g.min1, g.inv1, g.g1);
}
__global__ void __launch_bounds__(BE_THREADS) k_bin_event(const float *__restrict__ pos, int stride, int dim,
                                                          const int *__restrict__ ptr, float r, int phi_wrap, int cap_eff,
                                                          const int *__restrict__ bbox, int *__restrict__ cell_start,
                                                          float4 *__restrict__ tmp, float4 *__restrict__ pts,
                                                          int *__restrict__ perm, int *__restrict__ list,
                                                          int *__restrict__ count) {
    __shared__ int s_cnt[CMAX];          // histogram, then the scatter cursor
    __shared__ int s_start[CMAX + 1];    // event-local exclusive starts
    __shared__ int s_wtot[32];
    pdl_sync();
    const CellGrid g = make_grid(bbox, r, dim, phi_wrap);
    const int e = blockIdx.x, t = threadIdx.x, lane = t & 31, wp = t >> 5;
    const int lo = ptr[e], n = ptr[e + 1] - lo;
    s_cnt[t] = 0;
    __syncthreads();
    for (int i = t; i < n; i += BE_THREADS) {
        float a, b;
        atomicAdd(&s_cnt[cell_of(g, pos, stride, dim, (size_t)(lo + i), a, b)], 1);
    }
    __syncthreads();
    {   // exclusive scan of the 1024 counts (thread = cell)
        const int v = s_cnt[t];
        int tot;
        const int ex = warp_excl_scan_int(v, lane, tot);
        if (lane == 31) s_wtot[wp] = tot;
        __syncthreads();
        if (wp == 0) {
            int dummy;
            const int wex = warp_excl_scan_int(s_wtot[lane], lane, dummy);
            s_wtot[lane] = wex;
        }
        __syncthreads();
        const int st = s_wtot[wp] + ex;
        s_start[t] = st;
        if (t == CMAX - 1) s_start[CMAX] = st + v;
        cell_start[(size_t)e * (CMAX + 1) + t] = lo + st;
        if (t == CMAX - 1) cell_start[(size_t)e * (CMAX + 1) + CMAX] = lo + st + v;
        s_cnt[t] = 0;                                        // now the cursor
    }
    __syncthreads();
    for (int i = t; i < n; i += BE_THREADS) {
        float a, b;
        const int c = cell_of(g, pos, stride, dim, (size_t)(lo + i), a, b);
        const int k = atomicAdd(&s_cnt[c], 1);               // arbitrary order inside the cell: fixed by the ranking below
        tmp[lo + s_start[c] + k] = make_float4(a, b, __int_as_float(i), __int_as_float(c));
    }
    __syncthreads();                                         // this CTA's global writes are visible to this CTA
    for (int s0 = t & ~31; s0 < n; s0 += BE_THREADS) {       // warp-aligned: the ballot below
        const int s = s0 + lane;
        bool dense = false;
        if (s < n) {
            float4 p = __ldcg(&tmp[lo + s]);
            const int idx = __float_as_int(p.z), cell = __float_as_int(p.w);
            const int c0 = s_start[cell], c1 = s_start[cell + 1];
            int rank = 0;
            for (int k = c0; k < c1; ++k) rank += __float_as_int(__ldcg(&tmp[lo + k].z)) < idx;
            p.w = __int_as_float(cell | (e << 10));
            pts[lo + c0 + rank] = p;
            perm[lo + c0 + rank] = lo + idx;
            // population of the 3x3 neighbourhood (nbhd_runs with cpa = cpb = cp, on the shared starts)
            const int ce = cell / g.g1, cp = cell - ce * g.g1;
            int C = 0;
#pragma unroll
            for (int de = -1; de <= 1; ++de) {
                const int ce2 = ce + de;
                if (ce2 < 0 || ce2 >= g.g0) continue;
                const int *row = s_start + ce2 * g.g1;
                int l = cp - 1, h = cp + 1;
                if (phi_wrap) {
                    if (g.g1 >= 3 && 3 < g.g1) {
                        if (l < 0) { C += row[g.g1] - row[g.g1 - 1]; l = 0; }
                        else if (h >= g.g1) { C += row[1] - row[0]; h = g.g1 - 1; }
                    } else { l = 0; h = g.g1 - 1; }
                }
                l = l < 0 ? 0 : l;
                h = h >= g.g1 ? g.g1 - 1 : h;
                C += row[h + 1] - row[l];
            }
            dense = C > cap_eff;
        }
        const unsigned mk = __ballot_sync(0xffffffffu, dense);
        if (mk) {
            int base = 0;
            if (lane == 0) base = atomicAdd(count, __popc(mk));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (dense) list[base + __popc(mk & ((1u << lane) - 1u))] = lo + s;
        }
    }
}

// The counting-sort scatter (k_cell_fill) places the points of a cell in atomic (arbitrary) order.  Internal order
// must be reproducible run to run (it fixes every summation order downstream), so each point is re-ranked inside its
// cell by original index: thread per slot, O(cell population) compares against L1/L2-resident neighbours.
// Same launch, cap pre-filter: only a target whose 3x3 neighbourhood holds more than `cap` points can be capped at
// all -- a property of its CELL, hence of the slot (ranking permutes inside a cell).  Those slots (a few percent outside
// jet cores) are appended to a list for the warp-per-target kernel below; the order of the list is arbitrary and
// irrelevant (thr[s] depends on s alone).
__global__ void __launch_bounds__(256) k_cell_rank(const float4 *__restrict__ tmp, const int *__restrict__ ptr, int N,
                                                   int B, const int *__restrict__ cell_start,
                                                   float4 *__restrict__ pts, int *__restrict__ perm,
                                                   const GeoMeta *__restrict__ meta, const int *__restrict__ bbox,
                                                   int *__restrict__ list, int *__restrict__ count) {
    pdl_sync();
    const GeoMeta m = *meta;
    const CellGrid g = make_grid(bbox, m.r, m.dim, m.phi_wrap);
    const int lane = threadIdx.x & 31;
    for (int s0 = (blockIdx.x * blockDim.x + threadIdx.x) & ~31; s0 < N; s0 += gridDim.x * blockDim.x) {
        const int s = s0 + lane;
        bool dense = false;
        if (s < N) {
            const int e = find_event(ptr, B, s);
            float4 t = tmp[s];
            const int idx = __float_as_int(t.z), cell = __float_as_int(t.w);
            const int *cs = cell_start + (size_t)e * (CMAX + 1);
            const int lo = __ldg(cs + cell), hi = __ldg(cs + cell + 1);
            int rank = 0;
            for (int k = lo; k < hi; ++k) rank += __float_as_int(tmp[k].z) < idx;
            t.w = __int_as_float(cell | (e << 10));
            pts[lo + rank] = t;
            perm[lo + rank] = __ldg(ptr + e) + idx;
            const int ce = cell / g.g1, cp = cell - ce * g.g1;
            int rlo[6], rhi[6];
            const int nr = nbhd_runs(g, cs, ce, cp, cp, m.phi_wrap, rlo, rhi);
            int C = 0;
            for (int q = 0; q < nr; ++q) C += rhi[q] - rlo[q];
            dense = C > m.cap_eff;
        }
        const unsigned mk = __ballot_sync(0xffffffffu, dense);
        if (mk) {
            int base = 0;
            if (lane == 0) base = atomicAdd(count, __popc(mk));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (dense) list[base + __popc(mk & ((1u << lane) - 1u))] = s;
        }
    }
}

// thr[s] for the targets that can exceed the cap at all (3x3 neighbourhood population > cap): hits are marked in a
// per-warp shared bitmask indexed by the event-local ORIGINAL index, the cap-th set bit is the last admissible source.
__global__ void __launch_bounds__(256) k_geo_cap(const int *__restrict__ ptr, int N, int B, const GeoMeta *__restrict__ meta,
                                                 const int *__restrict__ bbox, const int *__restrict__ cell_start,
                                                 const float4 *__restrict__ pts, int mask_words, int *__restrict__ thr,
                                                 const int *__restrict__ list, const int *__restrict__ count,
                                                 int *__restrict__ stats) {
    pdl_sync();
    extern __shared__ unsigned int smask[];
    const GeoMeta m = *meta;
    const CellGrid g = make_grid(bbox, m.r, m.dim, m.phi_wrap);
    const int lane = threadIdx.x & 31;
    unsigned int *mask = smask + (size_t)(threadIdx.x >> 5) * mask_words;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    int local_max = 0, local_capped = 0;
    const int n_list = *count;
    for (int li = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; li < n_list; li += warps) {
        const int s = list[li];
        const float4 t = pts[s];
        const int e = pts_event(t.w), cell = pts_cell(t.w);
        const int base = __ldg(ptr + e), n_e = __ldg(ptr + e + 1) - base;
        const int ce = cell / g.g1, cp = cell - ce * g.g1;
        int rlo[6], rhi[6];
        const int nr = nbhd_runs(g, cell_start + (size_t)e * (CMAX + 1), ce, cp, cp, m.phi_wrap, rlo, rhi);
        int C = 0;
        for (int q = 0; q < nr; ++q) C += rhi[q] - rlo[q];
        if (C <= m.cap_eff) continue;                      // cannot be capped: thr stays INT_MAX-like
        const int words = (n_e + 31) >> 5;
        if (words > mask_words) {                          // host promised a smaller max_event_nodes
            if (lane == 0) atomicOr(stats + 3, GM_FLAG_EVENT_TOO_LARGE);
            continue;
        }
        int total = 0;                                     // pass 1: count only (most cap candidates stay below it)
        for (int q = 0; q < nr; ++q)
            for (int k = rlo[q] + lane; k < rhi[q]; k += 32) {
                const float4 c = pts[k];
                total += dist2(t.x, t.y, c.x, c.y, m.dim, m.phi_wrap) < m.r2;
            }
        total = warp_sum_int(total);
        local_max = max(local_max, total);
        if (total <= m.cap_eff) continue;
        for (int w = lane; w < words; w += 32) mask[w] = 0u;   // pass 2: mark hits by original index, find the cap-th
        __syncwarp();
        for (int q = 0; q < nr; ++q)
            for (int k = rlo[q] + lane; k < rhi[q]; k += 32) {
                const float4 c = pts[k];
                if (dist2(t.x, t.y, c.x, c.y, m.dim, m.phi_wrap) < m.r2) {
                    const int cl = __float_as_int(c.z);
                    atomicOr(mask + (cl >> 5), 1u << (cl & 31));
                }
            }
        __syncwarp();
        int found = 0x7fffffff, seen = 0;
        for (int w0 = 0; w0 < words; w0 += 32) {
            const unsigned int word = (w0 + lane < words) ? mask[w0 + lane] : 0u;
            int tot;
            const int before = seen + warp_excl_scan_int(__popc(word), lane, tot);
            if (before < m.cap_eff && before + __popc(word) >= m.cap_eff) {   // the cap-th hit (1-based) is in this word
                unsigned int v = word;
                for (int k = m.cap_eff - before - 1; k > 0; --k) v &= v - 1;
                found = ((w0 + lane) << 5) + __ffs(v) - 1;
            }
            seen += tot;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) found = min(found, __shfl_xor_sync(0xffffffffu, found, o));
        if (lane == 0) {
            local_max = max(local_max, seen);
            if (seen > m.cap_eff) { thr[s] = found; ++local_capped; }
        }
        __syncwarp();
    }
    if (lane == 0) {
        if (local_max) atomicMax(stats + 1, local_max);
        if (local_capped) { atomicAdd(stats + 2, local_capped); atomicOr(stats + 3, GM_FLAG_DEGREE_CAPPED); }
    }
}

namespace ag_big {
#include "geo_aggr.inc"
}  // namespace ag_big
#undef GEO_AG_CH
#undef GEO_AG_MINB
#ifndef GEO_AG_SMALL_CH
#define GEO_AG_SMALL_CH 320
#endif
#ifndef GEO_AG_SMALL_MINB
#define GEO_AG_SMALL_MINB 2
#endif
#define GEO_AG_CH GEO_AG_SMALL_CH
#define GEO_AG_MINB GEO_AG_SMALL_MINB
namespace ag_small {
#include "geo_aggr.inc"
}  // namespace ag_small

// ------------------------------------------------------------------------------------------ backward: route to sources
// S[j][c] = sum over targets i with arg(i,c) == j of d conv_i[c]   (max-aggregation sends each (target, channel)
// gradient to ONE edge; d conv = BatchNorm backward of G, evaluated on the fly).
// Work item = 32 consecutive SOURCES.  Every target that can point at them lies in their cell neighbourhood, whose
// arg / G / raw rows are streamed through shared memory by TMA (2-stage pipeline).  Each warp walks staged target
// rows; lane c looks at arg(i,c) and, if that source belongs to the item, adds d conv_i[c] into ITS column of a
// per-warp [32 sources][32 channels] accumulator — a lane owns its column, so there are no conflicts and no atomics.
// No distance test, no cap threshold, no transposed graph: arg == j already says the edge exists.  The four warp
// accumulators are folded in warp order and rows are walked in staged (= internal node) order: bit-deterministic.
constexpr int SC_CH = GEO_SC_CH;               // staged target rows per pipeline stage
#ifndef GEO_SC_STAGES
#define GEO_SC_STAGES 2
#endif
constexpr int SC_STAGES = GEO_SC_STAGES;
constexpr size_t SC_STAGE_BYTES = (size_t)3 * SC_CH * 128;                   // [arg | G | raw][SC_CH][32] x 4 B

// Warp-specialised like the forward kernel: a producer warp walks the items and segments and streams the three row
// arrays (chunk descriptor + TMA bulk copies, full / empty mbarriers, two stages), S2_CONSUMERS consumer warps walk the
// staged rows on their own -- the only consumer-wide barriers are the two around the fold of a segment's accumulators
// (the per-chunk CTA barriers of the first version were 3.1 stall cycles per issued instruction, ncu).
constexpr int S2_CONSUMERS = GEO_WARPS;
constexpr int S2_THREADS = (S2_CONSUMERS + 1) * 32;
constexpr size_t S2_STAGE_BYTES = SC_STAGE_BYTES + 128;                         // + chunk descriptor
constexpr size_t S2_ACC = SC_STAGES * S2_STAGE_BYTES;                           // [S2_CONSUMERS][32][32] f32
constexpr size_t S2_COEF = S2_ACC + (size_t)S2_CONSUMERS * 32 * 32 * 4;         // [5][32] f32
constexpr size_t S2_BAR = S2_COEF + 5 * 32 * 4;                                 // full[2], empty[2]
constexpr size_t S2_TOTAL = S2_BAR + 16 * SC_STAGES;
static_assert(S2_STAGE_BYTES % 128 == 0 && S2_BAR % 8 == 0, "scatter staging alignment");

__global__ void __launch_bounds__(S2_THREADS)
k_geo_scatter(const float *__restrict__ G, const float *__restrict__ raw, BnBwd bn, const int *__restrict__ arg,
              GeoView v, int N, float *__restrict__ S) {
    extern __shared__ __align__(128) unsigned char gs[];
    float(*s_acc)[32][32] = reinterpret_cast<float(*)[32][32]>(gs + S2_ACC);
    float *s_coef = reinterpret_cast<float *>(gs + S2_COEF);
    const uint32_t full0 = smem_u32(gs + S2_BAR), empty0 = full0 + 8u * SC_STAGES;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    const GeoMeta m = *v.meta;
    if (tid == 0) {
        for (int s = 0; s < SC_STAGES; ++s) { mbar_init(full0 + 8u * s, 1); mbar_init(empty0 + 8u * s, S2_CONSUMERS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    pdl_sync();                                          // the BatchNorm sums, G: the previous kernel's outputs
    bn_bwd_coeffs(bn, s_coef);
    if (warp < S2_CONSUMERS)
        for (int i = lane; i < 32 * 32; i += 32) (&s_acc[warp][0][0])[i] = 0.f;
    __syncthreads();

    const int n_items = (N + GEO_TPI - 1) / GEO_TPI;
    if (warp == S2_CONSUMERS) {
        // ================================================================== producer
        const CellGrid g = make_grid(v.bbox, m.r, m.dim, m.phi_wrap);
        uint32_t gc = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int n0 = item * GEO_TPI, icnt = min(GEO_TPI, N - n0);
            const bool valid = lane < icnt;
            const float pw = valid ? __ldg(&v.pts[n0 + lane].w) : 0.f;
            const int e = valid ? pts_event(pw) : -1;
            const int cell = pts_cell(pw);
            const int key = valid ? ((e << 5) | (cell / g.g1)) : -1;
            const int prev = __shfl_up_sync(0xffffffffu, key, 1);
            unsigned heads = __ballot_sync(0xffffffffu, valid && (lane == 0 || key != prev));
            while (heads) {
                const int a = __ffs(heads) - 1;
                heads &= heads - 1;
                const int b = heads ? __ffs(heads) - 1 : icnt;
                const int ev = __shfl_sync(0xffffffffu, e, a);
                const int ca = __shfl_sync(0xffffffffu, cell, a), cb = __shfl_sync(0xffffffffu, cell, b - 1);
                const int ce = ca / g.g1;
                int rlo[6], rhi[6], voff[7];
                nbhd_runs6(g, v.cell_start + (size_t)ev * (CMAX + 1), ce, ca - ce * g.g1, cb - ce * g.g1, m.phi_wrap, rlo, rhi);
                constexpr int nr = 6;                          // empty slots stay (zero rows): static indexing, no local memory
                int C = 0;
#pragma unroll
                for (int q = 0; q < 6; ++q) { voff[q] = C; C += rhi[q] - rlo[q]; }
                voff[6] = C;
                const int nck = (C + SC_CH - 1) / SC_CH;
                for (int k = 0; k < nck; ++k, ++gc) {
                    const int st = gc % SC_STAGES;
                    mbar_wait(empty0 + 8u * st, ((gc / SC_STAGES) & 1u) ^ 1u, 200);
                    if (lane == 0) {
                        unsigned char *stage = gs + st * S2_STAGE_BYTES;
                        const uint32_t s_dst = smem_u32(stage), bar = full0 + 8u * st;
                        const int c0 = k * SC_CH, c1 = min(C, c0 + SC_CH);
                        ChunkDesc *d = reinterpret_cast<ChunkDesc *>(stage + SC_STAGE_BYTES);
                        d->flags = (k == 0 ? AG_FIRST : 0) | (k == nck - 1 ? AG_LAST : 0);
                        d->n0 = n0; d->a = a; d->b = b; d->rows = c1 - c0;
                        mbar_expect_tx(bar, (uint32_t)(c1 - c0) * 384u);
#pragma unroll
                        for (int q = 0; q < 6; ++q) {
                            if (q >= nr) break;
                            const int lo = max(voff[q], c0), hi = min(voff[q + 1], c1);
                            if (lo >= hi) continue;
                            const size_t off = (size_t)(rlo[q] + (lo - voff[q])) * 128;
                            const uint32_t n = (uint32_t)(hi - lo) * 128u, row = (uint32_t)(lo - c0) * 128u;
                            bulk_g2s(s_dst + row, reinterpret_cast<const char *>(arg) + off, n, bar);
                            bulk_g2s(s_dst + SC_CH * 128u + row, reinterpret_cast<const char *>(G) + off, n, bar);
                            bulk_g2s(s_dst + 2u * SC_CH * 128u + row, reinterpret_cast<const char *>(raw) + off, n, bar);
                        }
                    }
                    __syncwarp();
                }
            }
        }
        const int st = gc % SC_STAGES;                      // end-of-work descriptor (no copies)
        mbar_wait(empty0 + 8u * st, ((gc / SC_STAGES) & 1u) ^ 1u);
        if (lane == 0) {
            reinterpret_cast<ChunkDesc *>(gs + st * S2_STAGE_BYTES + SC_STAGE_BYTES)->flags = AG_END;
            mbar_arrive(full0 + 8u * st);
        }
    } else {
        // ================================================================== consumers
        const float c_mean = s_coef[lane], c_inv = s_coef[H + lane], c_scale = s_coef[2 * H + lane];
        const float c_1 = s_coef[3 * H + lane], c_2 = s_coef[4 * H + lane];
        for (uint32_t gc = 0;; ++gc) {
            const int st = gc % SC_STAGES;
            const unsigned char *stage = gs + st * S2_STAGE_BYTES;
            mbar_wait(full0 + 8u * st, (gc / SC_STAGES) & 1u);
            const ChunkDesc *d = reinterpret_cast<const ChunkDesc *>(stage + SC_STAGE_BYTES);
            const int flags = d->flags;
            if (flags & AG_END) break;
            const int a = d->a, b = d->b, rows = d->rows, n0 = d->n0;
            const unsigned first = (unsigned)(n0 + a), span = (unsigned)(b - a);
            const int *s_arg = reinterpret_cast<const int *>(stage);
            const float *s_G = reinterpret_cast<const float *>(s_arg + SC_CH * 32);
            const float *s_R = s_G + SC_CH * 32;
            // 4 staged rows per trip: their arg / G / raw loads are issued together (unconditionally: shared memory,
            // no traffic to save), then the <= 4 read-modify-writes of the accumulator run in row order
            for (int r = warp; r < rows; r += 4 * S2_CONSUMERS) {
                unsigned rel[4];
                float dv[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int rr = min(r + u * S2_CONSUMERS, SC_CH - 1);
                    rel[u] = (r + u * S2_CONSUMERS < rows) ? (unsigned)s_arg[rr * 32 + lane] - first : 0xffffffffu;
                    const float xhat = (s_R[rr * 32 + lane] - c_mean) * c_inv;
                    dv[u] = c_scale * (s_G[rr * 32 + lane] - c_1 - xhat * c_2);
                }
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (rel[u] < span) s_acc[warp][a + rel[u]][lane] += dv[u];
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty0 + 8u * st);       // this warp is done with the stage
            if (flags & AG_LAST) {
                // fold the warp accumulators of this segment's sources in warp order, write S, clear for the next use
                asm volatile("bar.sync 1, %0;" ::"n"(S2_CONSUMERS * 32) : "memory");
                for (int idx = tid; idx < (b - a) * 32; idx += S2_CONSUMERS * 32) {
                    const int t = a + (idx >> 5), c = idx & 31;
                    float tot = 0.f;
#pragma unroll
                    for (int w = 0; w < S2_CONSUMERS; ++w) { tot += s_acc[w][t][c]; s_acc[w][t][c] = 0.f; }
                    S[(size_t)(n0 + t) * H + c] = tot;
                }
                asm volatile("bar.sync 1, %0;" ::"n"(S2_CONSUMERS * 32) : "memory");
            }
        }
    }
}

// BatchNorm moments of an [N, 32] activation, as a pass of its own: the forward geo kernel takes its items from a
// ticket counter (dense jet cells cost many times a forward-calorimeter cell: static schedules left the SMs idle
// 24-37 % of the kernel), so its per-CTA contents depend on the schedule; this kernel reads the freshly written,
// L2-resident rows with a STATIC assignment (warp w of the grid owns rows w, w + W, ...), which keeps the moments --
// and with them the whole training step -- bit-reproducible.  Lane = channel, 8 row loads in flight per lane,
// shifted sums per warp (shift = the warp's first row), Chan combination in warp / CTA order, last CTA finalises.
constexpr int RM_THREADS = 256;
__global__ void __launch_bounds__(RM_THREADS) k_row_moments(const float *__restrict__ x, int N, AggrTail tail) {
    pdl_sync();
    __shared__ float s_red[RM_THREADS / 8][3][H];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int W = (gridDim.x * RM_THREADS) >> 5;
    float cnt = 0.f, shift = 0.f, s1 = 0.f, s2 = 0.f;
    for (int r0 = blockIdx.x * (RM_THREADS / 32) + warp; r0 < N; r0 += 8 * W) {
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int r = r0 + u * W;
            v[u] = r < N ? __ldcg(x + (size_t)r * H + lane) : 0.f;
        }
        if (cnt == 0.f) shift = v[0];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            if (r0 + u * W < N) {
                const float d = v[u] - shift;
                cnt += 1.f; s1 += d; s2 = fmaf(d, d, s2);
            }
        }
    }
    const float rn = cnt > 0.f ? 1.f / cnt : 0.f;
    s_red[warp][0][lane] = cnt;
    s_red[warp][1][lane] = fmaf(s1, rn, shift);
    s_red[warp][2][lane] = fmaxf(s2 - s1 * s1 * rn, 0.f);
    __syncthreads();
    if (warp == 0) {
        Moments tot{0.f, 0.f, 0.f};
#pragma unroll
        for (int k = 0; k < RM_THREADS / 32; ++k)
            tot = moments_combine(tot, Moments{s_red[k][0][lane], s_red[k][1][lane], s_red[k][2][lane]});
        float *dst = tail.partials + (size_t)blockIdx.x * MOM_STRIDE;
        dst[lane] = tot.n; dst[H + lane] = tot.mean; dst[2 * H + lane] = tot.m2;
    }
    if (last_cta_done(tail.counter))
        bn_finalize_cta(tail.partials, gridDim.x, tail.bn, tail.momentum, tail.eps, tail.stat, s_red);
}

// ------------------------------------------------------------------------------------------ launchers
// One launch of k_geo_aggr<AGGR, ARG, Q4> of tuning namespace NS (a macro: the two namespaces hold same-named templates)
#define GEO_AGGR_LAUNCH(NS, ARG, Q4, SLOT)                                                                              \
    do {                                                                                                                \
        if (!occ[SLOT]) {                                                                                               \
            GM_CUDA_TRY(cudaFuncSetAttribute(NS::k_geo_aggr<AGGR, ARG, Q4>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                             (int)NS::AG_TOTAL));                                                       \
            occ[SLOT] = resident_ctas(NS::k_geo_aggr<AGGR, ARG, Q4>, AG_THREADS, NS::AG_TOTAL, 3);                      \
        }                                                                                                               \
        grid = persistent_grid(n_items, min(occ[SLOT], MAX_PARTIAL_CTAS / 148));                                        \
        split = tail.partials && (long long)n_items >= 8LL * grid;                                                      \
        const AggrTail in_kernel = split ? AggrTail{} : tail;                                                           \
        GM_LAUNCH_PDL((NS::k_geo_aggr<AGGR, ARG, Q4>), grid, AG_THREADS, NS::AG_TOTAL, st, Y, P, v, N, out, arg, in_kernel); \
    } while (0)

template <int AGGR>
int launch_geo_aggr(const float *Y, const float *P, const GeoView &v, int N, int B, float *out, int *arg,
                    const AggrTail &tail, cudaStream_t st) {
    static int occ[4] = {0, 0, 0, 0};
    const bool q4 = !arg && !tail.partials;              // inference: quarter-warp rows
    const int n_items = (N + GEO_TPI - 1) / GEO_TPI;
    int grid = 1;
    // Train mode, two ways to the BatchNorm moments: with >= 8 items per CTA the ticket schedule wins more than the
    // extra (L2-resident) pass of k_row_moments costs (c5: 233 -> 212 us); below that the static schedule with the
    // moments folded into the kernel is faster (c2: 72 vs 80 us).  Both are bit-reproducible.
    bool split = false;
    if (arg && AGGR == GM_AGGR_MAX && (long long)N < (long long)GEO_AG_SMALL_EVENT * B)
        GEO_AGGR_LAUNCH(ag_small, true, false, 3);       // training on DY/TT-sized events
    else if (arg) GEO_AGGR_LAUNCH(ag_big, true, false, 1);
    else if (q4) GEO_AGGR_LAUNCH(ag_big, false, true, 2);
    else GEO_AGGR_LAUNCH(ag_big, false, false, 0);
    GM_LAUNCH_CHECK();
    if (split) {
        const int mgrid = persistent_grid((N + 63) / 64, min(4, MAX_PARTIAL_CTAS / 148));
        GM_LAUNCH_PDL(k_row_moments, mgrid, RM_THREADS, 0, st, out, N, tail);
        GM_LAUNCH_CHECK();
    }
    return GM_OK;
}

}  // namespace

int geo_edge_aggr(int aggr, const float *Y, const float *P, const void *geo_ws, int N, int B, float *out, int *arg,
                  const AggrTail &tail, cudaStream_t st) {
    if (!geo_ws || B < 1) return GM_ERR_INVALID_ARGUMENT;
    const GeoView v = view_of(geo_ws, N, B);
    switch (aggr) {
        case GM_AGGR_MAX: return launch_geo_aggr<GM_AGGR_MAX>(Y, P, v, N, B, out, arg, tail, st);
        case GM_AGGR_ADD: return launch_geo_aggr<GM_AGGR_ADD>(Y, P, v, N, B, out, arg, tail, st);
        case GM_AGGR_MEAN: return launch_geo_aggr<GM_AGGR_MEAN>(Y, P, v, N, B, out, arg, tail, st);
    }
    return GM_ERR_INVALID_ARGUMENT;
}

// BatchNorm moments of a finished [N,32] layer output as a separate static-schedule pass (bit-reproducible): what the
// tcgen05 EdgeConv (geo_tc.cu, ticket-scheduled, no in-kernel moments) uses in training
int geo_row_moments(const float *out, int N, const AggrTail &tail, cudaStream_t st) {
    const int mgrid = persistent_grid((N + 63) / 64, min(4, MAX_PARTIAL_CTAS / 148));
    GM_LAUNCH_PDL(k_row_moments, mgrid, RM_THREADS, 0, st, out, N, tail);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

int geo_scatter_s(const float *G, const float *raw, const BnBwd &bn, const int *arg, const void *geo_ws, int N, int B,
                  float *S, cudaStream_t st) {
    if (!geo_ws || B < 1) return GM_ERR_INVALID_ARGUMENT;
    const GeoView v = view_of(geo_ws, N, B);
    static int occ = 0;
    if (!occ) {
        GM_CUDA_TRY(cudaFuncSetAttribute(k_geo_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S2_TOTAL));
        occ = resident_ctas(k_geo_scatter, S2_THREADS, S2_TOTAL, 3);
    }
    const int grid = persistent_grid((N + GEO_TPI - 1) / GEO_TPI, occ);
    GM_LAUNCH_PDL(k_geo_scatter, grid, S2_THREADS, S2_TOTAL, st, G, raw, bn, arg, v, N, S);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

}  // namespace gm

using namespace gm;

extern "C" {

size_t gm_geo_ws_bytes(int N, int B) { return geo_ws_layout(N < 0 ? 0 : N, B < 0 ? 0 : B, nullptr, nullptr); }

int gm_geo_build(const float *pos, int pos_stride, int dim, const int *ptr, int N, int B, double r,
                 int max_num_neighbors, int loop, int phi_wrap, int max_event_nodes, void *ws, size_t ws_bytes,
                 int *node_perm_out, int *stats_out, void *stream) {
    if (N < 0 || B < 1 || !ptr || !ws || (N > 0 && (!pos || !node_perm_out))) return GM_ERR_INVALID_ARGUMENT;
    if (!(r > 0.0) || max_num_neighbors < 1 || max_event_nodes < 0) return GM_ERR_INVALID_ARGUMENT;
    if (dim < 1 || dim > 2 || B >= GEO_MAX_EVENTS) return GM_ERR_UNSUPPORTED;
    if (ws_bytes < geo_ws_layout(N, B, nullptr, nullptr)) return GM_ERR_CAPACITY;
    cudaStream_t st = as_stream(stream);
    GeoWs w;
    geo_ws_layout(N, B, (char *)ws, &w);
    TimedRegion tr(GM_TIMER_RADIUS_GRAPH, st);
    const float rf = (float)r, r2 = (float)(r * r);          // r*r in double, then fp32
    const int cap_eff = loop ? max_num_neighbors : max_num_neighbors + 1;
    int *stats = stats_out ? stats_out : w.stats;             // the kernels count straight into the caller's words
    const GeoInit gi{w.meta, GeoMeta{rf, r2, dim, phi_wrap, loop, cap_eff, N, B}, w.ptr, w.sched, w.thr};
    // enough events of ordinary size to fill the GPU with one CTA each: k_bin_event (256 x 5k-candidate events: build
    // 278 -> 259 us); otherwise the four grid-wide kernels of bin_points (32 events: 32 CTAs are slower, 76 -> 94 us)
    const bool per_event = N > 0 && B >= GEO_BIN_EVENT_MIN_B && max_event_nodes > 0 && max_event_nodes <= GEO_BIN_EVENT_MAX;
    GM_TRY(bin_points(pos, pos_stride, dim, ptr, N, B, rf, phi_wrap, w.bbox, w.cell_cnt, w.cell_start, w.tmp, stats, st, &gi,
                      per_event));
    if (N > 0) {
        const int mask_words = (max_event_nodes + 31) / 32 + 1;
        int threads = 256;
        while (threads > 32 && (size_t)(threads / 32) * mask_words * 4 > 200 * 1024) threads >>= 1;
        const size_t smem = (size_t)(threads / 32) * mask_words * 4;
        if (smem > 200 * 1024) return GM_ERR_UNSUPPORTED;
        if (smem > 48 * 1024)
            GM_CUDA_TRY(cudaFuncSetAttribute(k_geo_cap, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        const int grid = persistent_grid(((long long)N * 32 + threads - 1) / threads, 2048 / threads);
        if (per_event)
            GM_LAUNCH_PDL(k_bin_event, B, BE_THREADS, 0, st, pos, pos_stride, dim, ptr, rf, phi_wrap, cap_eff, w.bbox,
                          w.cell_start, w.tmp, w.pts, node_perm_out, w.cap_list, w.sched + 2);
        else
            GM_LAUNCH_PDL(k_cell_rank, persistent_grid((N + 255) / 256, 8), 256, 0, st, w.tmp, ptr, N, B, w.cell_start, w.pts,
                          node_perm_out, w.meta, w.bbox, w.cap_list, w.sched + 2);
        GM_LAUNCH_CHECK();
        int *cap_list = w.cap_list;
        GM_LAUNCH_PDL(k_geo_cap, grid, threads, smem, st, ptr, N, B, w.meta, w.bbox, w.cell_start, w.pts, mask_words, w.thr,
                      cap_list, w.sched + 2, stats);
        GM_LAUNCH_CHECK();
    }
    return GM_OK;
}

}  // extern "C"
