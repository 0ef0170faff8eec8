// a7/a8/a9: per-event MET reduction + loss (model/net.py:49-62), resolution metric (model/net.py:92-157),
// fused flat AdamW (train.py:52,75), plus library-wide helpers and the generic edge_index -> CSR sort.
#include <vector>

#include "common.cuh"

namespace gm {

int g_last_cuda_error = 0;
long long g_launch_count = 0;
bool pdl_enabled() {
    static const bool on = [] { const char *e = getenv("GM_PDL"); return !(e && e[0] == '0'); }();
    return on;
}

namespace {
struct TimerSlot {
    std::vector<cudaEvent_t> beg, end;
    size_t used = 0;
};
TimerSlot g_timers[GM_NUM_TIMERS];
bool g_timing_on = false;
}  // namespace

TimedRegion::TimedRegion(int slot_, cudaStream_t st_) : slot(slot_), st(st_), on(g_timing_on) {
    if (!on) return;
    TimerSlot &t = g_timers[slot];
    if (t.used == t.beg.size()) {
        cudaEvent_t a, b;
        cudaEventCreate(&a);
        cudaEventCreate(&b);
        t.beg.push_back(a);
        t.end.push_back(b);
    }
    cudaEventRecord(t.beg[t.used], st);
}
TimedRegion::~TimedRegion() {
    if (!on) return;
    TimerSlot &t = g_timers[slot];
    cudaEventRecord(t.end[t.used], st);
    ++t.used;
}

int num_sms() {
    static int cached = 0;
    if (cached == 0) {
        int dev = 0, n = 0;
        if (cudaGetDevice(&dev) != cudaSuccess ||
            cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
            n = 148;  // B200
        cached = n;
    }
    return cached;
}

int exclusive_scan_inplace(int *data, int n, int *tile_totals, int *total_out, cudaStream_t st);

namespace {

constexpr int RED_THREADS = 256;

__device__ __forceinline__ float block_sum(float v, float *sm) {
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = 0.f;
    if (threadIdx.x == 0)
        for (int k = 0; k < (int)(blockDim.x >> 5); ++k) r += sm[k];
    __syncthreads();
    return r;  // valid in thread 0
}

// 1024 threads per event: at ~1.7k candidates per event the launch is one or two dependent load batches per thread
// (32 CTAs of 256 threads took 7 batches)
constexpr int MET_THREADS = 1024;
// one CTA per event: fixed thread-strided order + fixed tree => deterministic MET
__global__ void __launch_bounds__(MET_THREADS) k_met(const float *__restrict__ w, const float *__restrict__ pred,
                                                     int pstride, const int *__restrict__ ptr,
                                                     float *__restrict__ met) {
    pdl_sync();
    __shared__ float sm[MET_THREADS / 32];
    const int e = blockIdx.x;
    const int lo = ptr[e], hi = ptr[e + 1];
    float sx = 0.f, sy = 0.f;
    for (int n = lo + threadIdx.x; n < hi; n += MET_THREADS) {
        const float wn = w[n];
        const float *row = pred + (size_t)n * pstride;
        sx = fmaf(wn, row[0], sx);
        sy = fmaf(wn, row[1], sy);
    }
    sx = block_sum(sx, sm);
    sy = block_sum(sy, sm);
    if (threadIdx.x == 0) { met[2 * e] = sx; met[2 * e + 1] = sy; }
}

// MET + d loss / d w_n = ((METx+tx) px + (METy+ty) py) * grad_scale (1/B for loss_fn; 1 for the per-event SUM the
// data-parallel step all-reduces before dividing by the GLOBAL event count) in one launch (training): the CTA of an event has everything d loss / d w needs once its own
// MET is reduced; the second sweep over the event's rows hits L1/L2.  Same summation order as k_met.
__global__ void __launch_bounds__(MET_THREADS) k_met_grad(const float *__restrict__ w, const float *__restrict__ pred,
                                                          int pstride, const int *__restrict__ ptr,
                                                          const float *__restrict__ truth, int tstride, float grad_scale,
                                                          float *__restrict__ met, float *__restrict__ gw) {
    pdl_sync();
    __shared__ float sm[MET_THREADS / 32];
    __shared__ float s_g[2];
    const int e = blockIdx.x;
    const int lo = ptr[e], hi = ptr[e + 1];
    float sx = 0.f, sy = 0.f;
    for (int n = lo + threadIdx.x; n < hi; n += MET_THREADS) {
        const float wn = w[n];
        const float *row = pred + (size_t)n * pstride;
        sx = fmaf(wn, row[0], sx);
        sy = fmaf(wn, row[1], sy);
    }
    sx = block_sum(sx, sm);
    sy = block_sum(sy, sm);
    if (threadIdx.x == 0) {
        met[2 * e] = sx; met[2 * e + 1] = sy;
        s_g[0] = (sx + truth[(size_t)e * tstride]) * grad_scale;
        s_g[1] = (sy + truth[(size_t)e * tstride + 1]) * grad_scale;
    }
    __syncthreads();
    const float gx = s_g[0], gy = s_g[1];
    for (int n = lo + threadIdx.x; n < hi; n += MET_THREADS) {
        const float *row = pred + (size_t)n * pstride;
        gw[n] = fmaf(gx, row[0], gy * row[1]);
    }
}

__global__ void __launch_bounds__(RED_THREADS) k_loss(const float *__restrict__ met, const float *__restrict__ truth,
                                                      int tstride, int B, float *__restrict__ loss) {
    pdl_sync();
    __shared__ float sm[RED_THREADS / 32];
    float acc = 0.f;
    for (int e = threadIdx.x; e < B; e += RED_THREADS) {
        const float dx = met[2 * e] + truth[(size_t)e * tstride];
        const float dy = met[2 * e + 1] + truth[(size_t)e * tstride + 1];
        acc += fmaf(dx, dx, dy * dy);
    }
    acc = block_sum(acc, sm);
    if (threadIdx.x == 0) loss[0] = 0.5f * acc / (float)B;
}

__global__ void __launch_bounds__(RED_THREADS) k_segment_sum(const float *__restrict__ src, const int *__restrict__ ptr,
                                                             float *__restrict__ out) {
    __shared__ float sm[RED_THREADS / 32];
    const int e = blockIdx.x;
    float s = 0.f;
    for (int n = ptr[e] + threadIdx.x; n < ptr[e + 1]; n += RED_THREADS) s += src[n];
    s = block_sum(s, sm);
    if (threadIdx.x == 0) out[e] = s;
}

// model/net.py:138-145 `compute`: response, u_par, u_perp of a MET-like vector against qT = truth[:,0:2]
__global__ void __launch_bounds__(256) k_resolution(const float *__restrict__ met, const float *__restrict__ truth,
                                                    int tstride, int K, int B, float *__restrict__ out,
                                                    float *__restrict__ qt) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= B) return;
    const float *t = truth + (size_t)e * tstride;
    const float qx = t[0], qy = t[1];
    const float qq = fmaf(qx, qx, qy * qy);
    const float qn = sqrtf(qq);
    qt[e] = qn;
    for (int k = 0; k < K; ++k) {
        float vx, vy;
        if (k == 0) { vx = -met[2 * e]; vy = -met[2 * e + 1]; }
        else { vx = t[2 * k]; vy = t[2 * k + 1]; }
        const float resp = fmaf(vx, qx, vy * qy) / qq;
        const float px = resp * qx, py = resp * qy;
        const float u_par = sqrtf(fmaf(px, px, py * py)) - qn;
        const float ox = vx - px, oy = vy - py;
        const float u_perp = sqrtf(fmaf(ox, ox, oy * oy));
        out[((size_t)k * 3 + 0) * B + e] = u_perp;
        out[((size_t)k * 3 + 1) * B + e] = u_par;
        out[((size_t)k * 3 + 2) * B + e] = resp;
    }
}

__global__ void __launch_bounds__(256) k_adamw(float *__restrict__ p, const float *__restrict__ g,
                                               float *__restrict__ m, float *__restrict__ v, int n, float lr,
                                               float b1, float b2, float eps, float wd, float bc1, float bc2_sqrt,
                                               float gscale, const float *__restrict__ denom, StepStatus status) {
    pdl_wait();                                          // no early trigger: whatever follows may read weights up front
    if (denom) gscale /= fmaxf(denom[0], 1.f);
    if (blockIdx.x == 0 && threadIdx.x == 0) status.publish();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float gi = g[i] * gscale;
        float pi = p[i] * (1.f - lr * wd);
        const float mi = b1 * m[i] + (1.f - b1) * gi;
        const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
        const float denom = sqrtf(vi) / bc2_sqrt + eps;
        pi -= (lr / bc1) * (mi / denom);
        p[i] = pi; m[i] = mi; v[i] = vi;
    }
}

// ---------------------------------------------------------------- generic edge_index -> CSR
__global__ void __launch_bounds__(256) k_count_targets(const long long *__restrict__ ei, long long E, int N,
                                                       int *__restrict__ deg) {
    for (long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x; k < E; k += (long long)gridDim.x * blockDim.x) {
        long long t = ei[E + k];
        if (t >= 0 && t < N) atomicAdd(deg + t, 1);
    }
}
__global__ void __launch_bounds__(256) k_fill_unsorted(const long long *__restrict__ ei, long long E, int N,
                                                       const int *__restrict__ rowptr, int *__restrict__ cursor,
                                                       int *__restrict__ tmp) {
    for (long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x; k < E; k += (long long)gridDim.x * blockDim.x) {
        long long t = ei[E + k];
        if (t >= 0 && t < N) tmp[rowptr[t] + atomicAdd(cursor + t, 1)] = (int)ei[k];
    }
}
// warp per row: rank sort by source (ties are identical values, so their relative order is immaterial)
__global__ void __launch_bounds__(256) k_sort_rows(const int *__restrict__ rowptr, int N, const int *__restrict__ tmp,
                                                   int *__restrict__ col, const int *__restrict__ gate) {
    if (gate && *gate == 0) return;
    const int lane = threadIdx.x & 31;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    for (int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < N; i += warps) {
        const int beg = rowptr[i], d = rowptr[i + 1] - beg;
        for (int a = lane; a < d; a += 32) {
            const int va = tmp[beg + a];
            int rank = 0;
            for (int b = 0; b < d; ++b) {
                const int vb = tmp[beg + b];
                rank += (vb < va) || (vb == va && b < a);
            }
            col[beg + rank] = va;
        }
    }
}

// transposed CSR: histogram of sources, warp per target row
__global__ void __launch_bounds__(256) k_count_sources(const int *__restrict__ rowptr, const int *__restrict__ col, int N,
                                                       int *__restrict__ deg_t, const int *__restrict__ gate) {
    if (gate && *gate == 0) return;
    const int lane = threadIdx.x & 31;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    for (int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < N; i += warps)
        for (int e = rowptr[i] + lane; e < rowptr[i + 1]; e += 32) atomicAdd(deg_t + col[e], 1);
}
__global__ void __launch_bounds__(256) k_fill_transposed(const int *__restrict__ rowptr, const int *__restrict__ col, int N,
                                                         const int *__restrict__ rowptr_t, int *__restrict__ cursor,
                                                         int *__restrict__ tmp, const int *__restrict__ gate) {
    if (gate && *gate == 0) return;
    const int lane = threadIdx.x & 31;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    for (int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < N; i += warps)
        for (int e = rowptr[i] + lane; e < rowptr[i + 1]; e += 32) {
            const int j = col[e];
            tmp[rowptr_t[j] + atomicAdd(cursor + j, 1)] = i;
        }
}
__global__ void __launch_bounds__(256) k_scatter_add(const float *__restrict__ src, const long long *__restrict__ index,
                                                     long long n, float *__restrict__ out, long long dim_size) {
    for (long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x; k < n; k += (long long)gridDim.x * blockDim.x) {
        const long long t = index[k];
        if (t >= 0 && t < dim_size) atomicAdd(out + t, src[k]);
    }
}
__global__ void __launch_bounds__(256) k_index_stats(const long long *__restrict__ index, long long n,
                                                     long long *__restrict__ stats) {
    long long mx = -1;
    int unsorted = 0;
    for (long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x; k < n; k += (long long)gridDim.x * blockDim.x) {
        const long long v = index[k];
        mx = v > mx ? v : mx;
        if (k > 0 && index[k - 1] > v) unsorted = 1;
    }
    if (mx >= 0) atomicMax(stats, mx);
    if (unsorted) atomicExch((unsigned long long *)(stats + 1), 0ull);
}
__global__ void k_index_stats_init(long long *stats) { stats[0] = -1; stats[1] = 1; }

struct CsrWs { int *cursor; int *tmp; int *tiles; };
size_t csr_ws_layout(int N, long long E, char *base, CsrWs *w) {
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return base ? base + o : nullptr; };
    CsrWs t;
    t.cursor = (int *)take((size_t)(N + 1) * sizeof(int));
    t.tmp = (int *)take((size_t)(E > 0 ? E : 1) * sizeof(int));
    t.tiles = (int *)take(((size_t)(N + 1) / 2048 + 2) * sizeof(int));
    if (w) *w = t;
    return off;
}

}  // namespace
}  // namespace gm

using namespace gm;

extern "C" {

int gm_version(void) { return GM_VERSION; }
int gm_last_cuda_error(void) { return g_last_cuda_error; }
int gm_num_sms(void) { return num_sms(); }
long long gm_launch_count(int reset) {
    long long v = g_launch_count;
    if (reset) g_launch_count = 0;
    return v;
}
int gm_timing_enable(int on) {
    g_timing_on = on != 0;
    return GM_OK;
}
int gm_timing_read(int slot, float *total_ms_out, int *regions_out) {
    if (slot < 0 || slot >= GM_NUM_TIMERS || !total_ms_out || !regions_out) return GM_ERR_INVALID_ARGUMENT;
    TimerSlot &t = g_timers[slot];
    float total = 0.f;
    for (size_t k = 0; k < t.used; ++k) {
        GM_CUDA_TRY(cudaEventSynchronize(t.end[k]));
        float ms = 0.f;
        GM_CUDA_TRY(cudaEventElapsedTime(&ms, t.beg[k], t.end[k]));
        total += ms;
    }
    *total_ms_out = total;
    *regions_out = (int)t.used;
    t.used = 0;
    return GM_OK;
}

const char *gm_status_string(int status) {
    switch (status) {
        case GM_OK: return "ok";
        case GM_ERR_INVALID_ARGUMENT: return "invalid argument";
        case GM_ERR_CAPACITY: return "caller-provided buffer too small";
        case GM_ERR_UNSUPPORTED: return "unsupported shape or mode";
        case GM_ERR_CUDA: return "CUDA runtime error";
    }
    return "unknown status";
}

int gm_met_loss(const float *w, const float *pred, int pred_stride, const float *truth, int truth_stride,
                const int *ptr, int N, int B, float *met_out, float *loss_out, float *grad_w_out, void *stream) {
    return gm_met_loss_ex(w, pred, pred_stride, truth, truth_stride, ptr, N, B, B > 0 ? 1.0f / (float)B : 0.f, met_out,
                          loss_out, grad_w_out, stream);
}

int gm_met_loss_ex(const float *w, const float *pred, int pred_stride, const float *truth, int truth_stride,
                   const int *ptr, int N, int B, float grad_scale, float *met_out, float *loss_out, float *grad_w_out,
                   void *stream) {
    if (B <= 0 || N < 0 || !ptr || !truth || !met_out || !loss_out || (N > 0 && (!w || !pred)))
        return GM_ERR_INVALID_ARGUMENT;
    if (pred_stride < 2 || truth_stride < 2) return GM_ERR_INVALID_ARGUMENT;
    cudaStream_t st = as_stream(stream);
    if (grad_w_out && N > 0)
        GM_LAUNCH_PDL(k_met_grad, B, MET_THREADS, 0, st, w, pred, pred_stride, ptr, truth, truth_stride, grad_scale, met_out, grad_w_out);
    else
        GM_LAUNCH_PDL(k_met, B, MET_THREADS, 0, st, w, pred, pred_stride, ptr, met_out);
    GM_LAUNCH_CHECK();
    GM_LAUNCH_PDL(k_loss, 1, RED_THREADS, 0, st, met_out, truth, truth_stride, B, loss_out);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

int gm_segment_sum(const float *src, const int *ptr, int B, float *out, void *stream) {
    if (B < 0 || !ptr || !out) return GM_ERR_INVALID_ARGUMENT;
    if (B == 0) return GM_OK;
    k_segment_sum<<<B, RED_THREADS, 0, as_stream(stream)>>>(src, ptr, out);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

int gm_resolution(const float *met, const float *truth, int truth_stride, int truth_cols, int B, float *out,
                  float *qt_out, void *stream) {
    if (B < 0 || !met || !truth || !out || !qt_out || truth_cols < 6 || truth_stride < truth_cols)
        return GM_ERR_INVALID_ARGUMENT;
    if (B == 0) return GM_OK;
    const int K = truth_cols > 6 ? 5 : 3;
    if (K == 5 && truth_cols < 10) return GM_ERR_INVALID_ARGUMENT;
    k_resolution<<<(B + 255) / 256, 256, 0, as_stream(stream)>>>(met, truth, truth_stride, K, B, out, qt_out);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

int gm_adamw_step(float *params, const float *grads, float *exp_avg, float *exp_avg_sq, int n, float lr, float beta1,
                  float beta2, float eps, float weight_decay, int step, float grad_scale, void *stream) {
    return gm_adamw_step_ex(params, grads, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, weight_decay, step, grad_scale,
                            nullptr, nullptr, nullptr, nullptr, stream);
}

int gm_adamw_step_ex(float *params, const float *grads, float *exp_avg, float *exp_avg_sq, int n, float lr, float beta1,
                     float beta2, float eps, float weight_decay, int step, float grad_scale,
                     const float *grad_denominator_dev, const int *err_flags_dev, const int *graph_stats_dev,
                     int *status_host, void *stream) {
    if (n < 0 || step < 1 || (n > 0 && (!params || !grads || !exp_avg || !exp_avg_sq))) return GM_ERR_INVALID_ARGUMENT;
    if (n == 0) return GM_OK;
    const double bc1 = 1.0 - pow((double)beta1, (double)step);
    const double bc2 = 1.0 - pow((double)beta2, (double)step);
    GM_LAUNCH_PDL(k_adamw, persistent_grid((n + 255) / 256, 8), 256, 0, as_stream(stream), params, grads, exp_avg, exp_avg_sq, n,
                  lr, beta1, beta2, eps, weight_decay, (float)bc1, (float)sqrt(bc2), grad_scale, grad_denominator_dev,
                  StepStatus{err_flags_dev, graph_stats_dev, status_host, step});
    GM_LAUNCH_CHECK();
    return GM_OK;
}

size_t gm_csr_ws_bytes(int N, long long E) { return csr_ws_layout(N < 0 ? 0 : N, E < 0 ? 0 : E, nullptr, nullptr); }

int gm_edge_index_to_csr(const int64_t *edge_index, long long E, int N, void *ws, size_t ws_bytes, int *rowptr_out,
                         int *col_out, void *stream) {
    if (N < 0 || E < 0 || !ws || !rowptr_out || (E > 0 && (!edge_index || !col_out))) return GM_ERR_INVALID_ARGUMENT;
    if (ws_bytes < csr_ws_layout(N, E, nullptr, nullptr)) return GM_ERR_CAPACITY;
    cudaStream_t st = as_stream(stream);
    CsrWs w;
    csr_ws_layout(N, E, (char *)ws, &w);
    GM_CUDA_TRY(cudaMemsetAsync(rowptr_out, 0, (size_t)(N + 1) * sizeof(int), st));
    GM_CUDA_TRY(cudaMemsetAsync(w.cursor, 0, (size_t)(N + 1) * sizeof(int), st));
    if (E > 0) {
        const int g = persistent_grid((E + 255) / 256, 8);
        k_count_targets<<<g, 256, 0, st>>>((const long long *)edge_index, E, N, rowptr_out);
        GM_LAUNCH_CHECK();
        GM_TRY(exclusive_scan_inplace(rowptr_out, N + 1, w.tiles, nullptr, st));
        k_fill_unsorted<<<g, 256, 0, st>>>((const long long *)edge_index, E, N, rowptr_out, w.cursor, w.tmp);
        GM_LAUNCH_CHECK();
        k_sort_rows<<<persistent_grid(((long long)N * 32 + 255) / 256, 8), 256, 0, st>>>(rowptr_out, N, w.tmp, col_out, nullptr);
        GM_LAUNCH_CHECK();
    }
    return GM_OK;
}

int gm_csr_transpose(const int *rowptr, const int *col, int N, long long E, void *ws, size_t ws_bytes,
                     int *rowptr_t_out, int *col_t_out, const int *only_if_flag, void *stream) {
    if (N < 0 || E < 0 || !ws || !rowptr_t_out || (N > 0 && !rowptr) || (E > 0 && (!col || !col_t_out)))
        return GM_ERR_INVALID_ARGUMENT;
    if (ws_bytes < csr_ws_layout(N, E, nullptr, nullptr)) return GM_ERR_CAPACITY;
    cudaStream_t st = as_stream(stream);
    CsrWs w;
    csr_ws_layout(N, E, (char *)ws, &w);
    GM_CUDA_TRY(cudaMemsetAsync(rowptr_t_out, 0, (size_t)(N + 1) * sizeof(int), st));
    GM_CUDA_TRY(cudaMemsetAsync(w.cursor, 0, (size_t)(N + 1) * sizeof(int), st));
    if (N > 0 && E > 0) {
        const int g = persistent_grid(((long long)N * 32 + 255) / 256, 8);
        k_count_sources<<<g, 256, 0, st>>>(rowptr, col, N, rowptr_t_out, only_if_flag);
        GM_LAUNCH_CHECK();
        GM_TRY(exclusive_scan_inplace(rowptr_t_out, N + 1, w.tiles, nullptr, st));
        k_fill_transposed<<<g, 256, 0, st>>>(rowptr, col, N, rowptr_t_out, w.cursor, w.tmp, only_if_flag);
        GM_LAUNCH_CHECK();
        k_sort_rows<<<g, 256, 0, st>>>(rowptr_t_out, N, w.tmp, col_t_out, only_if_flag);
        GM_LAUNCH_CHECK();
    }
    return GM_OK;
}

int gm_scatter_add(const float *src, const int64_t *index, long long n, float *out, long long dim_size, void *stream) {
    if (n < 0 || dim_size < 0 || (n > 0 && (!src || !index || !out))) return GM_ERR_INVALID_ARGUMENT;
    if (n == 0) return GM_OK;
    k_scatter_add<<<persistent_grid((n + 255) / 256, 8), 256, 0, as_stream(stream)>>>(src, (const long long *)index, n,
                                                                                     out, dim_size);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

int gm_index_stats(const int64_t *index, long long n, int64_t *stats_out, void *stream) {
    if (n < 0 || !stats_out || (n > 0 && !index)) return GM_ERR_INVALID_ARGUMENT;
    k_index_stats_init<<<1, 1, 0, as_stream(stream)>>>((long long *)stats_out);
    GM_LAUNCH_CHECK();
    if (n > 0) {
        k_index_stats<<<persistent_grid((n + 255) / 256, 8), 256, 0, as_stream(stream)>>>((const long long *)index, n,
                                                                                         (long long *)stats_out);
        GM_LAUNCH_CHECK();
    }
    return GM_OK;
}

}  // extern "C"
