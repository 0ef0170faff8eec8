// Cell grid shared by the radius-graph kernels (radius_graph.cu) and the cell-staged network kernels (geo.cu):
// points of an event are binned into a uniform (eta, phi) grid whose cell edge is >= r, so every neighbour of a
// point lies in the 3x3 cells around it — in (event, cell)-sorted order that is <= 3 contiguous runs (6 with phi wrap).
#pragma once
#include "common.cuh"

namespace gm {

constexpr int GMAX = 32;                 // cells per dimension (<=)
constexpr int CMAX = GMAX * GMAX;        // cells per event (<=)
constexpr float CELL_SLACK = 1.0005f;    // cell edge >= r * slack  => hits are always in adjacent cells
constexpr float PI_F = 3.14159265358979323846f;
constexpr float TWO_PI_F = 6.28318530717958647692f;

__device__ __forceinline__ int float_to_ordered(float f) {
    int i = __float_as_int(f);
    return i >= 0 ? i : i ^ 0x7fffffff;
}
__device__ __forceinline__ float ordered_to_float(int i) {
    return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff);
}

struct CellGrid {
    float min0, min1, inv0, inv1;
    int g0, g1;
};

// Same deterministic function of (bbox, r) in every kernel.
__device__ __forceinline__ CellGrid make_grid(const int *__restrict__ bbox, float r, int dim, int phi_wrap) {
    CellGrid g;
    const float cs = r * CELL_SLACK;
    float lo0 = ordered_to_float(bbox[0]), hi0 = ordered_to_float(bbox[1]);
    float ext0 = hi0 - lo0;
    if (!(ext0 > 0.f) || !isfinite(ext0)) { lo0 = isfinite(lo0) ? lo0 : 0.f; ext0 = 0.f; }
    int n0 = (int)fminf((float)GMAX, floorf(ext0 / cs));
    g.g0 = n0 < 1 ? 1 : n0;
    g.min0 = lo0;
    g.inv0 = ext0 > 0.f ? (float)g.g0 / ext0 : 0.f;
    if (dim < 2) {
        g.g1 = 1; g.min1 = 0.f; g.inv1 = 0.f;
    } else if (phi_wrap) {
        int n1 = (int)fminf((float)GMAX, floorf(TWO_PI_F / cs));
        g.g1 = n1 < 1 ? 1 : n1;
        g.min1 = -PI_F;
        g.inv1 = (float)g.g1 / TWO_PI_F;
    } else {
        float lo1 = ordered_to_float(bbox[2]), hi1 = ordered_to_float(bbox[3]);
        float ext1 = hi1 - lo1;
        if (!(ext1 > 0.f) || !isfinite(ext1)) { lo1 = isfinite(lo1) ? lo1 : 0.f; ext1 = 0.f; }
        int n1 = (int)fminf((float)GMAX, floorf(ext1 / cs));
        g.g1 = n1 < 1 ? 1 : n1;
        g.min1 = lo1;
        g.inv1 = ext1 > 0.f ? (float)g.g1 / ext1 : 0.f;
    }
    return g;
}

__device__ __forceinline__ int cell_coord(float v, float lo, float inv, int g) {
    if (!isfinite(v)) return 0;
    int c = (int)floorf((v - lo) * inv);
    return c < 0 ? 0 : (c >= g ? g - 1 : c);
}

__device__ __forceinline__ int find_event(const int *__restrict__ ptr, int B, int node) {
    int lo = 0, hi = B;  // largest e with ptr[e] <= node
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (__ldg(ptr + mid) <= node) lo = mid; else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ float dist2(float e0, float p0, float e1, float p1, int dim, int phi_wrap) {
    // reference arithmetic: dist = fmaf(dphi, dphi, deta*deta), fp32
    float de = e1 - e0;
    float acc = __fmul_rn(de, de);
    if (dim >= 2) {
        float dp = p1 - p0;
        if (phi_wrap) {
            if (dp > PI_F) dp -= TWO_PI_F;
            else if (dp < -PI_F) dp += TWO_PI_F;
        }
        acc = __fmaf_rn(dp, dp, acc);
    }
    return acc;
}



// Contiguous runs [lo, hi) of the (event, cell)-sorted point array that cover every cell adjacent to the cells
// (ce, cpa..cpb) of one eta-row: the 3 eta-rows ce-1..ce+1, phi cells cpa-1..cpb+1 (periodic with phi_wrap).
// cs = cell_start of the event ([g0*g1 + 1] used).  No point is covered twice.  Returns the number of runs (<= 6).
__device__ __forceinline__ int nbhd_runs(const CellGrid &g, const int *__restrict__ cs, int ce, int cpa, int cpb,
                                         int phi_wrap, int (&rlo)[6], int (&rhi)[6]) {
    int n = 0;
#pragma unroll
    for (int de = -1; de <= 1; ++de) {
        const int ce2 = ce + de;
        if (ce2 < 0 || ce2 >= g.g0) continue;
        const int *row = cs + ce2 * g.g1;
        int lo = cpa - 1, hi = cpb + 1;
        if (phi_wrap) {
            if (g.g1 >= 3 && hi - lo + 1 < g.g1) {           // the wrapped cell is not already inside [lo, hi]
                if (lo < 0) { rlo[n] = __ldg(row + g.g1 - 1); rhi[n] = __ldg(row + g.g1); ++n; lo = 0; }
                else if (hi >= g.g1) { rlo[n] = __ldg(row); rhi[n] = __ldg(row + 1); ++n; hi = g.g1 - 1; }
            } else {                                          // the range (with wrap) covers the whole phi ring
                lo = 0; hi = g.g1 - 1;
            }
        }
        lo = lo < 0 ? 0 : lo;
        hi = hi >= g.g1 ? g.g1 - 1 : hi;
        rlo[n] = __ldg(row + lo); rhi[n] = __ldg(row + hi + 1); ++n;
    }
    return n;
}

// nbhd_runs for the staging producers: ALWAYS six slots (eta-row ce-1: wrap run, main run; ce: ...; ce+1: ...), an
// absent run is empty (lo == hi == 0), so the slots are statically indexed (registers, not a local-memory array) and
// the twelve cell_start loads are unconditional and independent of each other: one memory round trip per segment
// instead of up to twelve (the producer warp is what the consumers of a DY/TT-sized step wait for).  Same runs in
// the same order as nbhd_runs once the empty slots are dropped.
__device__ __forceinline__ void nbhd_runs6(const CellGrid &g, const int *__restrict__ cs, int ce, int cpa, int cpb,
                                           int phi_wrap, int (&rlo)[6], int (&rhi)[6]) {
#pragma unroll
    for (int de = -1; de <= 1; ++de) {
        const int ce2 = ce + de;
        const bool ok = ce2 >= 0 && ce2 < g.g0;
        const int *row = cs + (ok ? ce2 : ce) * g.g1;
        int lo = cpa - 1, hi = cpb + 1, wl = -1;
        if (phi_wrap) {
            if (g.g1 >= 3 && hi - lo + 1 < g.g1) {
                if (lo < 0) { wl = g.g1 - 1; lo = 0; }
                else if (hi >= g.g1) { wl = 0; hi = g.g1 - 1; }
            } else {
                lo = 0; hi = g.g1 - 1;
            }
        }
        lo = lo < 0 ? 0 : lo;
        hi = hi >= g.g1 ? g.g1 - 1 : hi;
        const int wc = wl < 0 ? 0 : wl;
        const int w0 = __ldg(row + wc), w1 = __ldg(row + wc + 1), m0 = __ldg(row + lo), m1 = __ldg(row + hi + 1);
        const int s = 2 * (de + 1);
        rlo[s] = (ok && wl >= 0) ? w0 : 0; rhi[s] = (ok && wl >= 0) ? w1 : 0;
        rlo[s + 1] = ok ? m0 : 0;          rhi[s + 1] = ok ? m1 : 0;
    }
}

}  // namespace gm
