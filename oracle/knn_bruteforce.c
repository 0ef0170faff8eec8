/*
 * ORACLE — TEST INFRASTRUCTURE ONLY.  Never linked into or called from the
 * product path (deepmetv2_b200/).  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may use it.
 *
 * CPU restatement of the k-nearest-neighbour graph the reference sketches at
 *   /root/reference/model/graph_met_network.py:63
 *     knn_graph(emb, k=20, batch=batch, loop=True)            (commented there)
 *   /root/reference/model/dynamic_reduction_network.py:86,95
 *     to_undirected(knn_graph(x, self.k, batch, loop=False, flow=...))
 *
 * The algorithm lives in the un-vendored third-party package torch_cluster
 * (≈1.5.9, README.md:9-17 of the reference); this file restates the published
 * semantics of its `knn_graph` (python wrapper) over its brute-force CUDA
 * `knn_kernel`:
 *   - knn_graph asks knn() for K = k (loop) or k + 1 (!loop) neighbours, swaps
 *     the rows for flow='source_to_target' and, when !loop, masks row != col
 *     AFTERWARDS — so a node keeps k + 1 neighbours when the self pair is not
 *     among its K best (more than k exact duplicates at lower indices);
 *   - per query n_y of event e, candidates n_x scan [ptr[e], ptr[e+1]) in
 *     ascending order; dist = sum_d (x[n_x,d] - y[n_y,d])^2 in fp32, dimension
 *     order; nvcc contracts upstream's `dist += df * df` to an FMA, so here
 *     dist = fmaf(df, df, dist) explicitly (as radius_bruteforce.c does), and
 *     the CUDA kernel uses __fsub_rn / __fmaf_rn to match;
 *   - a candidate replaces the current worst entry only when strictly closer,
 *     so ties go to the lower index; the K best leave ordered by (dist, index);
 *   - events with fewer than K nodes yield fewer neighbours (-1 entries are
 *     dropped upstream by the `col != -1` mask).
 * PARITY UNPINNED against upstream binaries (torch_cluster is not installable
 * here); pinned by the property tests in tests/test_oracle_knn.py (exact
 * integer coordinates against a stable sort).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

#ifdef __cplusplus
extern "C" {
#endif

/* nbr_out: [N, K] int64, -1 padded; deg_out: [N] int64.  Returns the number of edges, or -1. */
long long oracle_knn_graph(const float *x, int dim, const int64_t *ptr, int B, int k, int loop,
                           int64_t *nbr_out, int64_t *deg_out)
{
    const int K = k + (loop ? 0 : 1);
    if (k < 1 || dim < 1) return -1;
    float *bd = (float *)malloc(sizeof(float) * (size_t)K);
    int64_t *bi = (int64_t *)malloc(sizeof(int64_t) * (size_t)K);
    long long E = 0;
    for (int e = 0; e < B; ++e) {
        for (int64_t i = ptr[e]; i < ptr[e + 1]; ++i) {
            for (int s = 0; s < K; ++s) { bd[s] = INFINITY; bi[s] = -1; }
            const float *qi = x + (size_t)i * dim;
            for (int64_t j = ptr[e]; j < ptr[e + 1]; ++j) {
                const float *cj = x + (size_t)j * dim;
                float acc = 0.0f;
                for (int d = 0; d < dim; ++d) {
                    const float df = cj[d] - qi[d];
                    acc = fmaf(df, df, acc);
                }
                if (acc < bd[K - 1]) {
                    int s = K - 1;
                    while (s > 0 && bd[s - 1] > acc) { bd[s] = bd[s - 1]; bi[s] = bi[s - 1]; --s; }
                    bd[s] = acc; bi[s] = j;
                }
            }
            int64_t n = 0;
            for (int s = 0; s < K; ++s) {
                if (bi[s] < 0) continue;
                if (!loop && bi[s] == i) continue;
                nbr_out[(size_t)i * K + n++] = bi[s];
            }
            for (int64_t s = n; s < K; ++s) nbr_out[(size_t)i * K + s] = -1;
            deg_out[i] = n;
            E += n;
        }
    }
    free(bd); free(bi);
    return E;
}

#ifdef __cplusplus
}
#endif
