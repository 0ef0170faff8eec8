/*
 * ORACLE — TEST INFRASTRUCTURE ONLY.  Never linked into or called from the
 * product path (deepmetv2_b200/).  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may use it.
 *
 * CPU restatement of the fixed-radius neighbour search the reference calls at
 *   /root/reference/train.py:48, /root/reference/evaluate.py:88,
 *   /root/reference/plt_weight.py:122
 *     radius_graph(etaphi, r=0.4, batch, loop=True, max_num_neighbors=255)
 *
 * The algorithm itself lives in the un-vendored third-party package
 * torch_cluster (≈1.5.9, pinned only through README.md:9-17 of the reference);
 * this file restates the published semantics of its CUDA `radius_kernel`
 * (SURVEY.md §3.5):
 *   - per query node n_y of event e, scan n_x over [ptr[e], ptr[e+1]) in
 *     ascending index order;
 *   - dist = 0; for d in dims: dist += (x[n_x,d]-y[n_y,d])^2  — nvcc contracts
 *     this to fmaf, so for D=2: dist = fmaf(d1,d1, d0*d0) in fp32;
 *   - keep when dist < r2 (strict), r2 = (float)((double)r*(double)r);
 *   - stop after max_num_neighbors hits (=> the lowest source indices survive);
 *   - self pairs are kept when loop=True, dropped otherwise (radius_graph asks
 *     for max_num_neighbors+1 when loop=False and filters row==col afterwards;
 *     restated here as the same filter + cap).
 *   - edge_index[0] = source (n_x), edge_index[1] = target (n_y); output order
 *     is (target asc, source asc).
 * PARITY UNPINNED against upstream binaries (torch_cluster is not installable
 * here); pinned only by the property tests in tests/test_oracle_radius.py.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

#ifdef __cplusplus
extern "C" {
#endif

static inline float dist2_fma(const float *a, const float *b, int dim, int phi_wrap)
{
    /* d = 0 term: 0 + d0*d0 (fmaf(d0,d0,0) == round(d0*d0)). */
    float acc = 0.0f;
    for (int d = 0; d < dim; ++d) {
        float diff = a[d] - b[d];
        if (phi_wrap && d == 1) {
            /* opt-in periodic phi, restating data_dytt/generate_npz.py:26-27
               (dphi + pi) mod 2pi - pi, evaluated in fp32 */
            const float PI_F = 3.14159265358979323846f;
            const float TWO_PI_F = 6.28318530717958647692f;
            if (diff > PI_F) diff -= TWO_PI_F;
            else if (diff < -PI_F) diff += TWO_PI_F;
        }
        acc = fmaf(diff, diff, acc);
    }
    return acc;
}

/*
 * pos      [N, dim] fp32 row-major
 * ptr      [B+1]    int64 event boundaries (batch must be sorted)
 * Two-call protocol: pass col == NULL to only count (rowptr is filled).
 * Returns the number of edges, or -1 on bad arguments.
 */
long long oracle_radius_graph(const float *pos, int dim, const int64_t *ptr, int B,
                              double r, int max_num_neighbors, int loop, int phi_wrap,
                              int64_t *rowptr /* [N+1] */, int64_t *col /* [E] or NULL */)
{
    if (!pos || !ptr || !rowptr || dim < 1 || B < 0) return -1;
    const float r2 = (float)(r * r);
    long long e_total = 0;
    rowptr[0] = 0;
    for (int e = 0; e < B; ++e) {
        const int64_t lo = ptr[e], hi = ptr[e + 1];
        for (int64_t ny = lo; ny < hi; ++ny) {
            int count = 0;
            /* loop=False: upstream searches with max+1 and removes self afterwards */
            const int cap = loop ? max_num_neighbors : max_num_neighbors + 1;
            int emitted = 0;
            for (int64_t nx = lo; nx < hi; ++nx) {
                float d2 = dist2_fma(pos + nx * dim, pos + ny * dim, dim, phi_wrap);
                if (d2 < r2) {
                    if (loop || nx != ny) {
                        if (col) col[e_total + emitted] = nx;
                        ++emitted;
                    }
                    if (++count >= cap) break;
                }
            }
            e_total += emitted;
            rowptr[ny + 1] = e_total;
        }
    }
    return e_total;
}

#ifdef __cplusplus
}
#endif
