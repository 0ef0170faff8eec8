// Graph operators of the DynamicReductionNetwork variant -- CUDA forms of the upstream calls at
// /root/reference/model/dynamic_reduction_network.py:86-102, restated on the CPU in oracle/drn_oracle.py:
//   to_undirected(knn_graph(...))   -> gm_csr_to_undirected   (symmetrise + coalesce, sorted by (row, col))
//   normalized_cut_2d               -> gm_normalized_cut
//   graclus                         -> gm_graclus              (torch_cluster's CPU greedy matching, per event)
//   max_pool / max_pool_x           -> gm_cluster_max_pool     (consecutive cluster ids + per-cluster max)
// Every kernel takes the node count as min(N, *n_dev) when n_dev is given: the pooled graph's size only exists on the
// device and the whole DRN forward runs without a host synchronisation.  Integer work: results are bit-exact.
#include "common.cuh"

namespace gm {

int exclusive_scan_inplace(int *data, int n, int *tile_totals, int *total_out, cudaStream_t st);

namespace {

__device__ __forceinline__ int eff_n(int N, const int *n_dev) {
    if (!n_dev) return N;
    const int v = *n_dev;
    return v < 0 ? 0 : (v < N ? v : N);
}

// directed input: either a CSR (rowptr, col) or the padded neighbour table of the k-NN search (nbr [N,K], deg [N])
struct Directed {
    const int *rowptr, *col, *nbr, *deg;
    int K;
    __device__ __forceinline__ int count(int i) const { return nbr ? deg[i] : rowptr[i + 1] - rowptr[i]; }
    __device__ __forceinline__ int entry(int i, int s) const { return nbr ? nbr[(size_t)i * K + s] : col[rowptr[i] + s]; }
};

// rcnt[j] = number of rows that list j  (rcnt zeroed by the caller)
__global__ void __launch_bounds__(256) k_und_count(Directed g, int N, const int *__restrict__ n_dev, int *__restrict__ rcnt) {
    pdl_sync();
    const int n = eff_n(N, n_dev);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int c = g.count(i);
        for (int s = 0; s < c; ++s) {
            const int j = g.entry(i, s);
            if (j >= 0 && j < n) atomicAdd(rcnt + j, 1);
        }
    }
}

// capptr[i] = own(i) + rcnt[i]  (then scanned: row i of the scratch list holds own entries, then the reverse ones)
__global__ void __launch_bounds__(256) k_und_cap(Directed g, int N, const int *__restrict__ n_dev, const int *__restrict__ rcnt,
                                                 int *__restrict__ capptr) {
    pdl_sync();
    const int n = eff_n(N, n_dev);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i <= N; i += gridDim.x * blockDim.x)
        capptr[i] = i < n ? g.count(i) + rcnt[i] : 0;
}

__global__ void __launch_bounds__(256) k_und_fill(Directed g, int N, const int *__restrict__ n_dev, int *__restrict__ rcnt,
                                                  const int *__restrict__ capptr, int *__restrict__ tmp) {
    pdl_sync();
    const int n = eff_n(N, n_dev);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int c = g.count(i), base = capptr[i];
        for (int s = 0; s < c; ++s) {
            const int j = g.entry(i, s);
            tmp[base + s] = j;
            if (j >= 0 && j < n) {
                const int slot = atomicSub(rcnt + j, 1) - 1;     // any order: the row is ranked afterwards
                tmp[capptr[j] + g.count(j) + slot] = i;
            }
        }
    }
}

// warp per node: flag every entry that repeats an earlier one of the row (or is out of range); degU = #distinct
__global__ void __launch_bounds__(256) k_und_dedupe(int N, const int *__restrict__ n_dev, const int *__restrict__ capptr,
                                                    const int *__restrict__ tmp, unsigned char *__restrict__ dupf,
                                                    int *__restrict__ rowptr_out) {
    pdl_sync();
    const int n = eff_n(N, n_dev);
    const int lane = threadIdx.x & 31;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    for (int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i <= N; i += warps) {
        if (i >= n) {
            if (lane == 0) rowptr_out[i] = 0;
            continue;
        }
        const int a = capptr[i], len = capptr[i + 1] - a;
        int distinct = 0;
        for (int t = lane; t < len; t += 32) {
            const int v = tmp[a + t];
            bool dup = v < 0 || v >= n;
            for (int s = 0; s < t; ++s) dup |= tmp[a + s] == v;
            dupf[a + t] = dup ? 1 : 0;
            distinct += dup ? 0 : 1;
        }
        distinct = warp_sum_int(distinct);
        if (lane == 0) rowptr_out[i] = distinct;
    }
}

// warp per node: rank of every distinct entry among the row's distinct entries -> sorted, coalesced row
__global__ void __launch_bounds__(256) k_und_emit(int N, const int *__restrict__ n_dev, const int *__restrict__ capptr,
                                                  const int *__restrict__ tmp, const unsigned char *__restrict__ dupf,
                                                  const int *__restrict__ rowptr_out, int *__restrict__ col_out,
                                                  int *__restrict__ row_out, long long capacity,
                                                  long long *__restrict__ ei64) {
    pdl_sync();
    const int n = eff_n(N, n_dev);
    const int lane = threadIdx.x & 31;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    for (int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += warps) {
        const int a = capptr[i], len = capptr[i + 1] - a, o = rowptr_out[i];
        for (int t = lane; t < len; t += 32) {
            if (dupf[a + t]) continue;
            const int v = tmp[a + t];
            int rank = 0;
            for (int s = 0; s < len; ++s) rank += (!dupf[a + s] && tmp[a + s] < v) ? 1 : 0;
            const long long p = (long long)o + rank;
            if (p >= capacity) continue;
            if (col_out) col_out[p] = v;
            if (row_out) row_out[p] = i;
            if (ei64) { ei64[p] = i; ei64[capacity + p] = v; }
        }
    }
}

struct UndWs {
    int *rcnt, *capptr, *tmp, *tiles;
    unsigned char *dupf;
};
size_t und_layout(int N, long long Ed, char *base, UndWs *w) {
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; };
    const size_t nodes = (size_t)(N > 0 ? N : 1) + 1, ent = (size_t)(Ed > 0 ? Ed : 1) * 2;
    const size_t o_r = take(nodes * 4), o_c = take(nodes * 4), o_t = take(ent * 4), o_d = take(ent),
                 o_s = take((nodes / 1024 + 2) * 4);
    if (w) {
        w->rcnt = (int *)(base + o_r); w->capptr = (int *)(base + o_c); w->tmp = (int *)(base + o_t);
        w->dupf = (unsigned char *)(base + o_d); w->tiles = (int *)(base + o_s);
    }
    return off;
}

// ------------------------------------------------------------------------------------------ normalized cut
// 8 lanes per edge: ||x[row] - x[col]||_2 * (1/deg[row] + 1/deg[col]);  E = rowptr[N] on the device
__global__ void __launch_bounds__(256) k_normalized_cut(const float *__restrict__ x, int stride, int dim,
                                                        const int *__restrict__ rowptr, const int *__restrict__ col,
                                                        const int *__restrict__ row, int N, long long capacity,
                                                        float *__restrict__ w_out) {
    pdl_sync();
    long long E = rowptr[N];
    if (E > capacity) E = capacity;
    const int sub = threadIdx.x & 7;
    const long long groups = ((long long)gridDim.x * blockDim.x) >> 3;
    const long long Epad = (E + 3) / 4 * 4;      // whole warps stay converged for the shuffles
    for (long long p = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 3; p < Epad; p += groups) {
        const bool live = p < E;
        const int r = live ? row[p] : 0, c = live ? col[p] : 0;
        const float4 *xr = reinterpret_cast<const float4 *>(x + (size_t)r * stride);
        const float4 *xc = reinterpret_cast<const float4 *>(x + (size_t)c * stride);
        float acc = 0.f;
        for (int d4 = sub; d4 < dim / 4; d4 += 8) {
            const float4 a = xr[d4], b = xc[d4];
            float d;
            d = a.x - b.x; acc = fmaf(d, d, acc);
            d = a.y - b.y; acc = fmaf(d, d, acc);
            d = a.z - b.z; acc = fmaf(d, d, acc);
            d = a.w - b.w; acc = fmaf(d, d, acc);
        }
        acc += __shfl_xor_sync(0xffffffffu, acc, 4);
        acc += __shfl_xor_sync(0xffffffffu, acc, 2);
        acc += __shfl_xor_sync(0xffffffffu, acc, 1);
        if (live && sub == 0) {
            const float dr = 1.0f / (float)(rowptr[r + 1] - rowptr[r]);
            const float dc = 1.0f / (float)(rowptr[c + 1] - rowptr[c]);
            w_out[p] = sqrtf(acc) * (dr + dc);
        }
    }
}

// ------------------------------------------------------------------------------------------ graclus
// One warp per event (the matching is sequential in visiting order by definition; events are independent components).
// Per block of 32 visited nodes: (1) the block's rows are staged into shared memory with independent loads (the loads
// of the NEXT block are issued before the current block is committed, so their latency is hidden); (2) every lane
// picks the best unmatched neighbour of ITS node from the state at block start -- the sequential answer unless that
// neighbour gets matched by an earlier node of the same block (matching OTHER neighbours only removes candidates that
// were not the maximum); (3) a short commit loop walks the 32 nodes in order and re-evaluates just those.
constexpr int GR_BUF = 768;       // staged row entries per block of 32 visited nodes (per buffer)
__device__ __forceinline__ void cp_async4(void *smem_dst, const void *gmem_src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src)
                 : "memory");
}

struct GrBlock {                   // lane-held description of a block of 32 visited nodes
    int u, ra, len, off, total;
};

__device__ __forceinline__ GrBlock gr_describe(const int *__restrict__ rowptr, const int *__restrict__ visit, int base, int n,
                                               int n0, int lane) {
    GrBlock b;
    const int pos = n0 + lane;
    b.u = pos < n ? (visit ? visit[base + pos] : base + pos) : -1;
    b.ra = b.u >= 0 ? rowptr[b.u] : 0;
    b.len = b.u >= 0 ? rowptr[b.u + 1] - b.ra : 0;
    b.off = warp_excl_scan_int(b.len, lane, b.total);
    return b;
}

// entry t of the block -> global position in col / weight (search over the scanned row lengths)
__device__ __forceinline__ int gr_source(const GrBlock &b, int t) {
    int lo = 0;                                            // last q with off[q] <= t
#pragma unroll
    for (int step = 16; step > 0; step >>= 1) {
        const int probe = __shfl_sync(0xffffffffu, b.off, lo + step);
        if (probe <= t) lo += step;
    }
    const int a = __shfl_sync(0xffffffffu, b.ra, lo), o = __shfl_sync(0xffffffffu, b.off, lo);
    return a + t - o;
}

// best admissible entry of row [o, o + l) of the staged (or global) lists for node u: returns the neighbour or u
__device__ __forceinline__ int gr_choose(const int *__restrict__ cols, const float *__restrict__ ws, bool weighted, int o, int l,
                                         int u, int base, int n, const int *__restrict__ state) {
    unsigned bkey = 0u;
    int bv = u;
    for (int t = 0; t < l; ++t) {
        const int v = cols[o + t];
        const int lv = v - base;
        if (v == u || lv < 0 || lv >= n) continue;
        unsigned key;
        if (weighted) {
            const float w = ws[o + t];
            if (!(w >= 0.f)) continue;
            key = __float_as_uint(w + 0.f) + 1u;           // weights >= 0 order like unsigned integers
        } else {
            key = 1u;
        }
        if (state[lv] >= 0) continue;
        // weighted: largest weight, ties to the LATER entry; unweighted: the first unmatched entry
        if (weighted ? key >= bkey : bkey == 0u) { bkey = key; bv = v; }
    }
    return bv;
}

__global__ void __launch_bounds__(32) k_graclus(const int *__restrict__ rowptr, const int *__restrict__ col,
                                                const float *__restrict__ weight, const int *__restrict__ ptr, int B,
                                                const int *__restrict__ visit, int state_cap, int *__restrict__ cluster,
                                                int *__restrict__ partner) {
    extern __shared__ int s_gr[];
    __shared__ unsigned s_hit[32];
    int *s_col = s_gr;                                         // [2][GR_BUF]
    float *s_w = reinterpret_cast<float *>(s_gr + 2 * GR_BUF); // [2][GR_BUF]
    int *s_state = s_gr + 4 * GR_BUF;                          // [state_cap] cluster id or -1
    pdl_sync();
    const int lane = threadIdx.x;
    const bool weighted = weight != nullptr;
    for (int e = blockIdx.x; e < B; e += gridDim.x) {
        const int base = ptr[e], n = ptr[e + 1] - base;
        if (n <= 0) continue;
        // the matched state: shared memory when the event fits, else the output array itself (slow, still exact)
        int *state = n <= state_cap ? s_state : cluster + base;
        for (int l = lane; l < n; l += 32) state[l] = -1;
        __syncwarp();
        GrBlock cur = gr_describe(rowptr, visit, base, n, 0, lane);
        // The rows of the NEXT block travel global -> shared memory as 4-byte cp.async copies into the other half of a
        // double buffer while the current block is chosen and committed.  (Register-held prefetches completed one by
        // one: ncu showed a full memory round trip on each of a block's ~11 load pairs, 45 % of the kernel.)
        auto stage_async = [&](const GrBlock &b, int half) {
            if (b.total <= GR_BUF) {
                const int first = __shfl_sync(0xffffffffu, b.ra, 0);
                for (int t0 = 0; t0 < b.total; t0 += 32) {             // warp-uniform trip count (gr_source shuffles)
                    const int t = t0 + lane;
                    // ascending visiting order: the rows of 32 consecutive nodes are one contiguous range of the CSR
                    const int src = visit ? gr_source(b, min(t, b.total - 1)) : first + t;
                    if (t < b.total) {
                        cp_async4(s_col + half * GR_BUF + t, col + src);
                        if (weighted) cp_async4(s_w + half * GR_BUF + t, weight + src);
                    }
                }
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
        int half = 0;
        stage_async(cur, half);
        for (int n0 = 0; n0 < n; n0 += 32) {
            const int cnt = min(32, n - n0);
            const bool staged = cur.total <= GR_BUF;
            asm volatile("cp.async.wait_all;" ::: "memory");
            __syncwarp();
            const int *cols = staged ? s_col + half * GR_BUF : col;
            const float *ws = staged ? s_w + half * GR_BUF : weight;
            const int row_o = staged ? cur.off : cur.ra;
            // (2) speculative choice from the state at block start, every lane for its own node
            int vsel = cur.u;
            if (cur.u >= 0 && state[cur.u - base] < 0)
                vsel = gr_choose(cols, ws, weighted, row_o, cur.len, cur.u, base, n, state);
            // next block: describe it and put its loads in flight before the sequential part
            GrBlock nxt = cur;
            if (n0 + 32 < n) {
                nxt = gr_describe(rowptr, visit, base, n, n0 + 32, lane);
                stage_async(nxt, half ^ 1);
            }
            // (3a) commit in visiting order, first WITHOUT touching the state: a speculative choice stands unless an
            // earlier node of this block took one of its two ends.  Lane q knows which lanes could (hitU: their choice
            // is my node; hitV: their choice or their node is my choice); walking q = 0 .. cnt-1 with the running mask
            // of committed lanes is one shuffle per node instead of two state loads, a store and a warp barrier.  The
            // walk stops at the first node whose chosen neighbour was taken (it has to choose again from the real
            // state); everything before it is written out in parallel and (3b) continues from there.
            int q0 = 0;
            if (!visit && state == s_state) {
                const int u = cur.u;
                const bool act = u >= 0 && state[u - base] < 0;
                const bool pair = act && vsel != u;
                const int vj = pair ? vsel - base - n0 : -1;               // my choice as a lane of this block, if it is one
                s_hit[lane] = 0u;
                __syncwarp();
                if (pair && vj >= 0 && vj < 32) atomicOr(&s_hit[vj], 1u << lane);
                __syncwarp();
                const unsigned hitU = s_hit[lane];
                unsigned hitV = 0u;
                const unsigned pairs = __ballot_sync(0xffffffffu, pair);
                if (pair) {
                    hitV = __match_any_sync(pairs, vsel) & ~(1u << lane);     // same choice
                    if (vj >= 0 && vj < 32) hitV |= 1u << vj;                 // my choice is a node of this block
                }
                unsigned done = 0u;
                int stop = cnt;
                for (int q = 0; q < cnt; ++q) {
                    int code = 0;                                             // 0 nothing, 1 commit, 2 must choose again
                    if (act && !(hitU & done)) code = (pair && (hitV & done)) ? 2 : 1;
                    code = __shfl_sync(0xffffffffu, code, q);
                    if (code == 2) { stop = q; break; }
                    if (code == 1) done |= 1u << q;
                }
                if (done & (1u << lane)) {
                    const int c = min(u, vsel);
                    state[u - base] = c;
                    state[vsel - base] = c;
                }
                __syncwarp();
                q0 = stop;
            }
            // (3b) the rest of the block against the real state
            for (int q = q0; q < cnt; ++q) {
                const int u = __shfl_sync(0xffffffffu, cur.u, q);
                int v = __shfl_sync(0xffffffffu, vsel, q);
                if (state[u - base] >= 0) continue;                               // matched earlier in this block
                if (v != u && state[v - base] >= 0) {                             // the chosen neighbour was taken: redo
                    const int o = __shfl_sync(0xffffffffu, row_o, q), l = __shfl_sync(0xffffffffu, cur.len, q);
                    v = gr_choose(cols, ws, weighted, o, l, u, base, n, state);   // every lane computes the same answer
                }
                const int c = min(u, v);
                if (lane == 0) { state[u - base] = c; state[v - base] = c; }      // shared memory only: no global store
                __syncwarp();                                                     // sits between two dependent steps
            }
            __syncwarp();
            cur = nxt;
            half ^= 1;
        }
        // write-out: cluster ids, and the partner derived from them (the non-representative member names both)
        for (int l = lane; l < n; l += 32) {
            if (state != cluster + base) cluster[base + l] = state[l];
            if (partner) partner[base + l] = -1;
        }
        __syncwarp();
        if (partner) {
            for (int l = lane; l < n; l += 32) {
                const int c = state[l];
                if (c != base + l) { partner[base + l] = c; partner[c] = base + l; }
            }
        }
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------------ pooling
__global__ void __launch_bounds__(256) k_pool_flags(const int *__restrict__ cluster, int N, const int *__restrict__ n_dev,
                                                    int *__restrict__ scan) {
    pdl_sync();
    const int n = eff_n(N, n_dev);
    for (int u = blockIdx.x * blockDim.x + threadIdx.x; u <= N; u += gridDim.x * blockDim.x)
        scan[u] = (u < n && cluster[u] == u) ? 1 : 0;
}

// 16 lanes per node (float4 each at dim 64): x_out[newid] = max(x[u], x[partner])
__global__ void __launch_bounds__(256) k_pool_apply(const float *__restrict__ x, int dim, const int *__restrict__ cluster,
                                                    const int *__restrict__ partner, const int *__restrict__ ptr, int N, int B,
                                                    const int *__restrict__ n_dev, const int *__restrict__ scan,
                                                    float *__restrict__ x_out, int *__restrict__ ptr_out,
                                                    int *__restrict__ newid_out) {
    pdl_sync();
    const int n = eff_n(N, n_dev);
    const int q = dim / 4;
    const long long total = (long long)n * q;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        const int u = (int)(t / q), d4 = (int)(t - (long long)u * q);
        const int cu = cluster[u];
        if (d4 == 0 && newid_out) newid_out[u] = scan[cu];
        if (cu != u) continue;
        float4 a = reinterpret_cast<const float4 *>(x + (size_t)u * dim)[d4];
        const int v = partner[u];
        if (v >= 0) {
            const float4 b = reinterpret_cast<const float4 *>(x + (size_t)v * dim)[d4];
            a.x = fmaxf(a.x, b.x); a.y = fmaxf(a.y, b.y); a.z = fmaxf(a.z, b.z); a.w = fmaxf(a.w, b.w);
        }
        reinterpret_cast<float4 *>(x_out + (size_t)scan[u] * dim)[d4] = a;
    }
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e <= B; e += gridDim.x * blockDim.x) {
        const int s = min(ptr[e], n);
        ptr_out[e] = scan[s];
    }
}

}  // namespace

// internal entry points shared with drn_net.cu
int drn_to_undirected(const int *rowptr, const int *col, const int *nbr, const int *deg, int K, int N, const int *n_dev,
                      void *ws, size_t ws_bytes, int *rowptr_out, int *col_out, int *row_out, long long capacity,
                      long long *ei64, cudaStream_t st) {
    const long long Ed = nbr ? (long long)N * K : (capacity + 1) / 2;     // CSR input: capacity >= 2 x directed edges
    UndWs w;
    if (ws_bytes < und_layout(N, Ed, nullptr, nullptr)) return GM_ERR_CAPACITY;
    und_layout(N, Ed, (char *)ws, &w);
    Directed g{rowptr, col, nbr, deg, K};
    GM_CUDA_TRY(cudaMemsetAsync(w.rcnt, 0, ((size_t)N + 1) * sizeof(int), st));
    const int gn = persistent_grid(((long long)N + 256) / 256, 8);
    const int gw = persistent_grid(((long long)N * 32 + 255) / 256 + 1, 8);
    GM_LAUNCH_PDL(k_und_count, gn, 256, 0, st, g, N, n_dev, w.rcnt);
    GM_LAUNCH_CHECK();
    GM_LAUNCH_PDL(k_und_cap, gn, 256, 0, st, g, N, n_dev, (const int *)w.rcnt, w.capptr);
    GM_LAUNCH_CHECK();
    GM_TRY(exclusive_scan_inplace(w.capptr, N + 1, w.tiles, nullptr, st));
    GM_LAUNCH_PDL(k_und_fill, gn, 256, 0, st, g, N, n_dev, w.rcnt, (const int *)w.capptr, w.tmp);
    GM_LAUNCH_CHECK();
    GM_LAUNCH_PDL(k_und_dedupe, gw, 256, 0, st, N, n_dev, (const int *)w.capptr, (const int *)w.tmp, w.dupf, rowptr_out);
    GM_LAUNCH_CHECK();
    GM_TRY(exclusive_scan_inplace(rowptr_out, N + 1, w.tiles, nullptr, st));
    GM_LAUNCH_PDL(k_und_emit, gw, 256, 0, st, N, n_dev, (const int *)w.capptr, (const int *)w.tmp,
                  (const unsigned char *)w.dupf, (const int *)rowptr_out, col_out, row_out, capacity, ei64);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

int drn_normalized_cut(const float *x, int stride, int dim, const int *rowptr, const int *col, const int *row, int N,
                       long long capacity, float *w_out, cudaStream_t st) {
    const int g = persistent_grid((capacity * 8 + 255) / 256, 8);
    GM_LAUNCH_PDL(k_normalized_cut, g, 256, 0, st, x, stride, dim, rowptr, col, row, N, capacity, w_out);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

int drn_graclus(const int *rowptr, const int *col, const float *weight, const int *ptr, int N, int B, int men,
                const int *visit, int *cluster, int *partner, cudaStream_t st) {
    (void)N;
    int cap = men > 0 ? men : 1;
    const int max_cap = (200 * 1024 - 4 * GR_BUF * 4) / 4;
    if (cap > max_cap) cap = max_cap;            // larger events run with the state in global memory
    const size_t smem = (size_t)(4 * GR_BUF + cap) * 4;
    static bool attr_set = false;
    if (!attr_set) {
        GM_CUDA_TRY(cudaFuncSetAttribute(k_graclus, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_set = true;
    }
    GM_LAUNCH_PDL(k_graclus, B, 32, smem, st, rowptr, col, weight, ptr, B, visit, cap, cluster, partner);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

int drn_max_pool(const float *x, int dim, const int *cluster, const int *partner, const int *ptr, int N, int B,
                 const int *n_dev, int *scan, int *tiles, float *x_out, int *ptr_out, int *newid_out, cudaStream_t st) {
    const int gn = persistent_grid(((long long)N + 256) / 256, 8);
    GM_LAUNCH_PDL(k_pool_flags, gn, 256, 0, st, cluster, N, n_dev, scan);
    GM_LAUNCH_CHECK();
    GM_TRY(exclusive_scan_inplace(scan, N + 1, tiles, nullptr, st));
    const int ga = persistent_grid(((long long)N * (dim / 4) + 255) / 256 + 1, 8);
    GM_LAUNCH_PDL(k_pool_apply, ga, 256, 0, st, x, dim, cluster, partner, ptr, N, B, n_dev, (const int *)scan, x_out,
                  ptr_out, newid_out);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

}  // namespace gm

using namespace gm;

extern "C" {

size_t gm_undirected_ws_bytes(int N, long long E_directed_capacity) {
    return und_layout(N < 0 ? 0 : N, E_directed_capacity, nullptr, nullptr);
}

int gm_csr_to_undirected(const int *rowptr, const int *col, const int *nbr, const int *deg, int K, int N,
                         const int *n_dev, void *ws, size_t ws_bytes, int *rowptr_out, int *col_out, int *row_out,
                         long long capacity, int64_t *edge_index64_out, void *stream) {
    if (N < 0 || !ws || !rowptr_out || capacity < 0) return GM_ERR_INVALID_ARGUMENT;
    const bool table = nbr != nullptr;
    if (N > 0 && (table ? (!deg || K < 1) : (!rowptr || !col))) return GM_ERR_INVALID_ARGUMENT;
    cudaStream_t st = as_stream(stream);
    if (N == 0) {
        GM_CUDA_TRY(cudaMemsetAsync(rowptr_out, 0, sizeof(int), st));
        return GM_OK;
    }
    return drn_to_undirected(rowptr, col, nbr, deg, K, N, n_dev, ws, ws_bytes, rowptr_out, col_out, row_out, capacity,
                             (long long *)edge_index64_out, st);
}

int gm_normalized_cut(const float *x, int x_stride, int dim, const int *rowptr, const int *col, const int *row,
                      int N, long long capacity, float *weight_out, void *stream) {
    if (N < 0 || capacity < 0 || (N > 0 && (!x || !rowptr || !col || !row || !weight_out))) return GM_ERR_INVALID_ARGUMENT;
    if (dim < 4 || (dim & 3) || (x_stride & 3) || (reinterpret_cast<uintptr_t>(x) & 15)) return GM_ERR_UNSUPPORTED;
    if (N == 0 || capacity == 0) return GM_OK;
    return drn_normalized_cut(x, x_stride, dim, rowptr, col, row, N, capacity, weight_out, as_stream(stream));
}

int gm_graclus(const int *rowptr, const int *col, const float *weight, const int *ptr, int N, int B,
               int max_event_nodes, const int *visit, int *cluster_out, int *partner_out, void *stream) {
    if (N < 0 || B < 0 || (N > 0 && (!rowptr || !col || !ptr || !cluster_out || B < 1))) return GM_ERR_INVALID_ARGUMENT;
    if (N == 0) return GM_OK;
    return drn_graclus(rowptr, col, weight, ptr, N, B, max_event_nodes, visit, cluster_out, partner_out, as_stream(stream));
}

int gm_cluster_max_pool(const float *x, int dim, const int *cluster, const int *partner, const int *ptr, int N,
                        int B, const int *n_dev, void *scan_ws, size_t scan_ws_bytes, float *x_out, int *ptr_out,
                        int *newid_out, void *stream) {
    if (N < 0 || B < 0 || !scan_ws || !ptr_out || (N > 0 && (!x || !cluster || !partner || !ptr || !x_out)))
        return GM_ERR_INVALID_ARGUMENT;
    if (dim < 4 || (dim & 3)) return GM_ERR_UNSUPPORTED;
    UndWs w;
    if (scan_ws_bytes < und_layout(N, 0, nullptr, nullptr)) return GM_ERR_CAPACITY;
    und_layout(N, 0, (char *)scan_ws, &w);
    cudaStream_t st = as_stream(stream);
    if (N == 0) {
        GM_CUDA_TRY(cudaMemsetAsync(ptr_out, 0, ((size_t)B + 1) * sizeof(int), st));
        return GM_OK;
    }
    return drn_max_pool(x, dim, cluster, partner, ptr, N, B, n_dev, w.capptr, w.tiles, x_out, ptr_out, newid_out, st);
}

}  // extern "C"
