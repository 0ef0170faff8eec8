// Workspace of the cell-staged ("geo") path: the radius graph kept IMPLICIT.
//
// gm_geo_build bins the points of every event into the (eta, phi) cell grid of cells.cuh and leaves, in this
// workspace, the points in INTERNAL node order = sorted by (event, cell, original index), the per-event cell
// starts, and one int per node (`thr`) that carries the max_num_neighbors cap.  An edge j -> i of the reference's
// radius_graph (train.py:48) exists iff
//     same event  &&  dist2(i, j) < r2  &&  local_index(j) <= thr[i]  &&  (loop || j != i)
// because torch_cluster keeps the max_num_neighbors LOWEST source indices of a target: thr[i] is the local index of
// the last admissible source (INT_MAX when the target is not capped).  The EdgeConv kernels of geo.cu re-evaluate
// this predicate on cell neighbourhoods staged in shared memory instead of reading an edge list.
#pragma once
#include "cells.cuh"

namespace gm {

struct GeoMeta {          // written once per build by a device kernel; read by every geo kernel
    float r, r2;
    int dim, phi_wrap, loop, cap_eff;   // cap_eff = max_num_neighbors (+1 when !loop: torch_cluster's radius_graph)
    int N, B;
};

struct GeoWs {
    GeoMeta *meta;
    int *bbox;            // [4] ordered-int min0, max0, min1, max1
    int *ptr;             // [B+1] copy of the caller's event boundaries (the workspace must outlive the caller's tensor)
    int *cell_cnt;        // [B*CMAX]   (left at zero after a build)
    int *cell_start;      // [B*(CMAX+1)] start of every cell in internal order (absolute node index)
    float4 *tmp;          // [N] points grouped by (event, cell), arbitrary order inside a cell
    float4 *pts;          // [N] (eta, phi, local original index bits, event << 10 | cell) in internal order
    int *thr;             // [N] internal order
    int *cap_list;        // [N] slots whose 3x3 neighbourhood can exceed the cap (first sched[2] entries)
    int *stats;           // [4] E (-1: not counted), max degree seen in dense cells, #capped, flags
    int *sched;           // [4] ticket / finished-CTA counters of the dynamically scheduled (eval) forward kernel,
                          //     [2] = length of the cap-candidate list
};

// pts[].w carries the node's cell (10 bits: CMAX = 1024) and its event, so the staging warps get both from the one
// load they do anyway instead of a 6-deep dependent binary search over ptr[] per work item (B < 2^21)
static_assert(CMAX == 1024, "pts.w packing");
constexpr int GEO_MAX_EVENTS = 1 << 21;
__device__ __forceinline__ int pts_cell(float w) { return __float_as_int(w) & (CMAX - 1); }
__device__ __forceinline__ int pts_event(float w) { return (int)((unsigned)__float_as_int(w) >> 10); }

inline size_t geo_ws_layout(int N, int B, char *base, GeoWs *w) {
    size_t off = 0;
    auto take = [&](size_t bytes) {
        size_t o = off;
        off = align_up(off + bytes, 256);
        return base ? base + o : nullptr;
    };
    const size_t nodes = (size_t)(N > 0 ? N : 1), events = (size_t)(B > 0 ? B : 1);
    GeoWs t;
    t.meta = (GeoMeta *)take(sizeof(GeoMeta));
    t.bbox = (int *)take(4 * sizeof(int));
    t.stats = (int *)take(4 * sizeof(int));
    t.sched = (int *)take(4 * sizeof(int));
    t.ptr = (int *)take((events + 1) * sizeof(int));
    t.cell_cnt = (int *)take(events * CMAX * sizeof(int));
    t.cell_start = (int *)take(events * (CMAX + 1) * sizeof(int));
    t.tmp = (float4 *)take(nodes * sizeof(float4));
    t.pts = (float4 *)take(nodes * sizeof(float4));
    t.thr = (int *)take(nodes * sizeof(int));
    t.cap_list = (int *)take(nodes * sizeof(int));
    if (w) *w = t;
    return off;
}

// What gm_geo_build has to reset per build (meta, ticket counters, its copy of ptr, the "not capped" default of thr):
// rides along in the first kernel of bin_points (one launch less on a latency-bound chain).  meta == nullptr: nothing.
struct GeoInit {
    GeoMeta *meta;
    GeoMeta value;
    int *ptr_copy, *sched, *thr;
};
__device__ __forceinline__ void geo_init_body(const GeoInit &gi, const int *__restrict__ ptr, int *stats) {
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x, gsize = gridDim.x * blockDim.x;
    if (gtid == 0) {
        *gi.meta = gi.value;
        stats[0] = -1;                                   // E is never counted on this path
        gi.sched[0] = 0; gi.sched[1] = 0; gi.sched[2] = 0; gi.sched[3] = 0;
    }
    for (int i = gtid; i <= gi.value.B; i += gsize) gi.ptr_copy[i] = ptr[i];
    for (int i = gtid; i < gi.value.N; i += gsize) gi.thr[i] = 0x7f7f7f7f;      // "not capped"
}

// implemented in radius_graph.cu: bbox -> cell histogram -> per-event scan -> scatter of (eta, phi, idx, cell);
// prologue_only: reset + bbox only (the caller bins the cells itself)
int bin_points(const float *pos, int pos_stride, int dim, const int *ptr, int N, int B, float r, int phi_wrap,
               int *bbox, int *cell_cnt, int *cell_start, float4 *sorted, int *stats, cudaStream_t st,
               const GeoInit *geo_init = nullptr, bool prologue_only = false);

// implemented in geo.cu, used by net_forward.cu / net_backward.cu
struct AggrTail;
struct BnBwd;
int geo_edge_aggr(int aggr, const float *Y, const float *P, const void *geo_ws, int N, int B, float *out, int *arg,
                  const AggrTail &tail, cudaStream_t st);
int geo_row_moments(const float *out, int N, const AggrTail &tail, cudaStream_t st);
// implemented in geo_tc.cu: the same layer as a per-edge tcgen05 contraction (tf32 / bf16); arg (optional, max only) =
// arg-max source node per (target, channel) for the backward
int geo_edgeconv_tc(int aggr, int precision, const float *X, const float *P, const float *W, const void *geo_ws, int N,
                    int B, float *out, int *arg, cudaStream_t st);
int geo_scatter_s(const float *G, const float *raw, const BnBwd &bn, const int *arg, const void *geo_ws, int N, int B,
                  float *S, cudaStream_t st);

}  // namespace gm
