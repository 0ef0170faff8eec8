// Staging machinery shared by the cell-staged EdgeConv kernels (geo.cu: factored fp32; geo_tc.cu: per-edge tcgen05):
// mbarrier / TMA bulk-copy PTX, the chunk descriptor, the stage layout, the producer warp that streams the cell
// neighbourhoods of 32-node work items through a 2-stage ring, and the per-target hit-list builder.
#pragma once
#include "geo_layout.cuh"
#include "net_layout.cuh"

namespace gm {
namespace {

constexpr int GEO_TPI = 32;          // nodes per work item (one warp-width of consecutive internal nodes)
constexpr int AG_STAGES = 2;
constexpr unsigned GEO_SPIN_LIMIT = 1u << 20;

// ------------------------------------------------------------------------------------------ PTX: mbarrier + TMA bulk copy
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// potentially blocking: the thread is suspended in hardware until the phase completes or ~hint_ns elapse, so waiting
// warps do not compete for issue slots with the warps that still work
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity, uint32_t hint_ns) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity), "r"(hint_ns) : "memory");
    return ok != 0;
}
// bounded (a bug must trap, not hang); failed polls back off with nanosleep so that waiting warps do not take issue
// slots from the warps that still work (ncu: 14 % of the executed instructions were try_wait polls without it)
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, unsigned sleep_ns = 128) {
    if (mbar_try_wait(bar, parity, 1000u)) return;
#pragma unroll 1
    for (unsigned it = 0; it < GEO_SPIN_LIMIT; ++it) {
        __nanosleep(sleep_ns);
        if (mbar_try_wait(bar, parity, 1000u)) return;
    }
    __trap();
}
// global -> shared bulk copy through the TMA unit; bytes % 16 == 0, both addresses 16-byte aligned
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

struct GeoView {       // device pointers of a built GeoWs (passed by value)
    const GeoMeta *meta;
    const int *bbox, *cell_start, *ptr, *thr;
    const float4 *pts;
    int *sched;        // [0..1] dynamic-schedule ticket + finished-CTA count (self-resetting; zeroed by gm_geo_build)
};

inline GeoView view_of(const void *geo_ws, int N, int B) {
    GeoWs w;
    geo_ws_layout(N, B, (char *)geo_ws, &w);
    return GeoView{w.meta, w.bbox, w.cell_start, w.ptr, w.thr, w.pts, w.sched};
}

// A work item = GEO_TPI consecutive internal nodes.  It is cut into SEGMENTS of equal (event, eta-row); the
// neighbourhood of a segment whose nodes sit in phi cells cpa..cpb is the <= 6 runs of nbhd_runs, flattened into
// "virtual rows" 0..C-1 and staged CH rows at a time.
struct ChunkDesc {     // written by the producer into the stage it fills, read by the consumers after the full barrier
    int flags;         // AG_FIRST / AG_LAST chunk of a segment, AG_END of the CTA's work
    int n0;            // first node of the item
    int a, b;          // targets [a, b) of the item form the segment
    int rows;          // staged rows of this chunk
    int vbase;         // virtual row of staged row 0
    int nr;            // runs of the segment's neighbourhood
    int rlo[6];        // first internal node of run q
    int voff[7];       // first virtual row of run q; voff[nr] = C
};
constexpr int AG_FIRST = 1, AG_LAST = 2, AG_END = 4;
static_assert(sizeof(ChunkDesc) <= 128, "descriptor slot");

// One pipeline stage: { (CH + 1) rows x 128 B (last row: identity), CH points x 16 B, the item's 32 target points
// x 16 B, their 32 cap thresholds, the chunk descriptor, [TROWS: the item's 32 target rows x 128 B, first chunk of a
// segment only] }
template <int CH, bool TROWS>
struct StageL {
    static constexpr int ROWS = CH;
    static constexpr size_t OFF_PTS = (size_t)(CH + 1) * 128;
    static constexpr size_t OFF_TGT = OFF_PTS + (size_t)CH * 16;
    static constexpr size_t OFF_THR = OFF_TGT + (size_t)GEO_TPI * 16;
    static constexpr size_t OFF_DESC = OFF_THR + (size_t)GEO_TPI * 4;
    static constexpr size_t OFF_TROWS = OFF_DESC + 128;
    static constexpr size_t BYTES = OFF_TROWS + (TROWS ? (size_t)GEO_TPI * 128 : 0);
    static_assert(BYTES % 128 == 0, "stage alignment");
};

__device__ __forceinline__ int node_of_row(const ChunkDesc *d, int row) {
    const int vr = d->vbase + row;
    int q = 0;
    while (q + 1 < d->nr && vr >= d->voff[q + 1]) ++q;
    return d->rlo[q] + (vr - d->voff[q]);
}

// The producer warp: walks the CTA's items (static: item = blockIdx + k * gridDim; dynamic: tickets from v.sched), cuts
// them into segments, computes the neighbourhood runs and streams points + 128-byte rows of `rows_src` through the ring
// with TMA bulk copies, SL::ROWS rows at a time; every chunk carries its descriptor and the item's target points / cap
// thresholds (and, with TROWS, the item's own rows of `rows_src` on the first chunk of a segment).
//   full[s]  : 1 arrival + TMA bytes;   empty[s] : one arrival per consumer warp.
template <class SL, bool TROWS, int NS = AG_STAGES>
__device__ __forceinline__ void geo_produce(unsigned char *gs, uint32_t full0, uint32_t empty0, const GeoView &v,
                                            const GeoMeta &m, const float *__restrict__ rows_src, int N, bool dynamic,
                                            int lane, unsigned sleep_ns = 400) {
    const CellGrid g = make_grid(v.bbox, m.r, m.dim, m.phi_wrap);
    const int n_items = (N + GEO_TPI - 1) / GEO_TPI;
    uint32_t gc = 0;                                    // chunks issued so far; chunk gc lives in stage gc % NS
    int item = blockIdx.x;
    for (;;) {
        if (dynamic) {
            if (lane == 0) item = atomicAdd(v.sched, 1);
            item = __shfl_sync(0xffffffffu, item, 0);
        }
        if (item >= n_items) break;
        const int n0 = item * GEO_TPI, icnt = min(GEO_TPI, N - n0);
        const bool valid = lane < icnt;
        const float pw = valid ? __ldg(&v.pts[n0 + lane].w) : 0.f;         // event << 10 | cell
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
            int rlo[6], rhi[6], voff[7];                // every lane computes the (uniform) runs: broadcast loads
            nbhd_runs6(g, v.cell_start + (size_t)ev * (CMAX + 1), ce, ca - ce * g.g1, cb - ce * g.g1, m.phi_wrap, rlo, rhi);
            constexpr int nr = 6;                          // empty slots stay (zero rows): static indexing, no local memory
            int C = 0;
#pragma unroll
            for (int q = 0; q < 6; ++q) { voff[q] = C; C += rhi[q] - rlo[q]; }
            voff[6] = C;
            const int nck = (C + SL::ROWS - 1) / SL::ROWS;
            const int crow = min(SL::ROWS, (((C + nck - 1) / nck) + 31) & ~31);    // even split, multiple of 32
            for (int k = 0; k < nck; ++k, ++gc) {
                const int st = gc % NS;
                mbar_wait(empty0 + 8u * st, ((gc / NS) & 1u) ^ 1u, sleep_ns);   // stage consumed (or never used)
                if (lane == 0) {
                    unsigned char *stage = gs + st * SL::BYTES;
                    const uint32_t s_stage = smem_u32(stage), bar = full0 + 8u * st;
                    const int c0 = k * crow, c1 = min(C, c0 + crow);
                    ChunkDesc *d = reinterpret_cast<ChunkDesc *>(stage + SL::OFF_DESC);
                    d->flags = (k == 0 ? AG_FIRST : 0) | (k == nck - 1 ? AG_LAST : 0);
                    d->n0 = n0; d->a = a; d->b = b; d->rows = c1 - c0; d->vbase = c0; d->nr = nr;
#pragma unroll
                    for (int q = 0; q < 6; ++q) { d->rlo[q] = rlo[q]; d->voff[q] = voff[q]; }
                    d->voff[6] = voff[6];
                    const bool trows = TROWS && k == 0;
                    mbar_expect_tx(bar, (uint32_t)(c1 - c0) * (128u + 16u) + (uint32_t)icnt * 16u + 128u +
                                            (trows ? (uint32_t)icnt * 128u : 0u));
                    bulk_g2s(s_stage + (uint32_t)SL::OFF_TGT, v.pts + n0, (uint32_t)icnt * 16u, bar);
                    bulk_g2s(s_stage + (uint32_t)SL::OFF_THR, v.thr + n0, 128u, bar);
                    if (trows)
                        bulk_g2s(s_stage + (uint32_t)SL::OFF_TROWS, rows_src + (size_t)n0 * H, (uint32_t)icnt * 128u, bar);
#pragma unroll
                    for (int q = 0; q < 6; ++q) {
                        if (q >= nr) break;
                        const int lo = max(voff[q], c0), hi = min(voff[q + 1], c1);
                        if (lo >= hi) continue;
                        const size_t node = (size_t)(rlo[q] + (lo - voff[q]));
                        const uint32_t n = (uint32_t)(hi - lo), row = (uint32_t)(lo - c0);
                        bulk_g2s(s_stage + (uint32_t)SL::OFF_PTS + row * 16u, v.pts + node, n * 16u, bar);
                        bulk_g2s(s_stage + row * 128u, rows_src + node * H, n * 128u, bar);
                    }
                }
                __syncwarp();
            }
        }
        item += gridDim.x;
    }
    const int st = gc % NS;                             // end-of-work descriptor (no copies)
    mbar_wait(empty0 + 8u * st, ((gc / NS) & 1u) ^ 1u);
    if (lane == 0) {
        reinterpret_cast<ChunkDesc *>(gs + st * SL::BYTES + SL::OFF_DESC)->flags = AG_END;
        mbar_arrive(full0 + 8u * st);
    }
}

// Hit list of one target against the staged chunk: BYTE OFFSETS (row * 128) of the rows whose point passes the edge
// predicate of geo_layout.cuh, in staged order, padded with 16 x the identity row (row CH).  th = cap threshold of the
// target, excl = local index to exclude (the target itself when loop == 0, -1 otherwise).  Returns the hit count.
template <int CH>
__device__ __forceinline__ int build_hit_list(const float4 *s_pts, int rows, float px, float py, int th, int excl,
                                              const GeoMeta &m, int lane, unsigned *list) {
    int nh = 0;
    const unsigned lt = (1u << lane) - 1u;
    // 128 candidates per trip, their 4 point loads issued up front: the list stores below may alias the points as
    // far as the compiler knows, so it cannot hoist the next load over them itself.  Reads past `rows` stay inside
    // the stage (CH + 32 <= points + target area) and are masked by c < rows.
    for (int base = 0; base < rows; base += 128) {
        float4 q[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) q[u] = s_pts[base + 32 * u + lane];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int c = base + 32 * u + lane;
            const int cl = __float_as_int(q[u].z);
            const bool hit = (dist2(px, py, q[u].x, q[u].y, m.dim, m.phi_wrap) < m.r2) & (cl <= th) & (cl != excl) & (c < rows);
            const unsigned mk = __ballot_sync(0xffffffffu, hit);
            unsigned *slot = list + nh + __popc(mk & lt);   // address formed unconditionally => predicated store
            if (hit) *slot = (unsigned)c << 7;
            nh += __popc(mk);
        }
    }
    if (lane < 16) list[nh + lane] = (unsigned)CH << 7;      // list capacity: CH + 16
    __syncwarp();
    return nh;
}

}  // namespace
}  // namespace gm
