// Row f4: the per-epoch qT-binned response / resolution summary of /root/reference/evaluate.py:110-156 on the device.
// The reference concatenates every batch's numpy arrays on the host, digitizes qT into 10-GeV bins and calls
// np.mean / np.quantile per (MET flavour, bin).  Here the per-event arrays of an evaluation epoch stay in device
// memory ([K,3,ld]: u_perp, u_par, response per flavour, as gm_resolution writes them, and qT [M]) and ONE launch
// produces, per flavour and bin, mean(R), (q84 - q16)/2 of u_perp and u_par, and the same for u / mean(R).
// One CTA per (bin, flavour): order statistics by an exact 4-pass radix select over the bin's members (no sort, no
// scratch), np.quantile's default linear interpolation, sums in double.  Empty bins yield NaN (np.mean of an empty slice).
#include "common.cuh"

namespace gm {
namespace {

constexpr int RS_THREADS = 256;

__device__ __forceinline__ unsigned orderable(float v) {          // monotone float -> unsigned
    const unsigned u = __float_as_uint(v);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float from_orderable(unsigned k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}
// np.digitize(x, arange(0, n_edges * width, width)): number of edges <= x
__device__ __forceinline__ int digitize(float x, float width, int n_edges) {
    if (!(x >= 0.f)) return 0;
    int c = (int)(x / width);
    if (c > n_edges - 1) c = n_edges - 1;
    while (c + 1 <= n_edges - 1 && (float)(c + 1) * width <= x) ++c;
    while (c > 0 && (float)c * width > x) --c;
    return c + 1;
}

// k-th smallest (0-based) of the members' values: exact radix select, 8 bits per pass
__device__ unsigned select_kth(const float *__restrict__ v, const float *__restrict__ qt, int M, int bin, float width,
                               int n_edges, unsigned k, unsigned *s_hist, unsigned *s_pick) {
    unsigned prefix = 0, mask = 0;
    for (int shift = 24; shift >= 0; shift -= 8) {
        for (int t = threadIdx.x; t < 256; t += RS_THREADS) s_hist[t] = 0;
        __syncthreads();
        for (int i = threadIdx.x; i < M; i += RS_THREADS) {
            if (digitize(qt[i], width, n_edges) != bin) continue;
            const unsigned key = orderable(v[i]);
            if ((key & mask) == prefix) atomicAdd(&s_hist[(key >> shift) & 0xffu], 1u);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned acc = 0, b = 0;
            for (; b < 256; ++b) {
                if (acc + s_hist[b] > k) break;
                acc += s_hist[b];
            }
            s_pick[0] = b; s_pick[1] = k - acc;
        }
        __syncthreads();
        prefix |= s_pick[0] << shift;
        mask |= 0xffu << shift;
        k = s_pick[1];
        __syncthreads();
    }
    return prefix;
}

// np.quantile(a, q) (method 'linear') of the bin's members
__device__ float quantile(const float *__restrict__ v, const float *__restrict__ qt, int M, int bin, float width, int n_edges,
                          int n, double q, unsigned *s_hist, unsigned *s_pick) {
    const double pos = (double)(n - 1) * q;
    const int lo = (int)floor(pos), hi = min(lo + 1, n - 1);
    const double g = pos - (double)lo;
    const float a = from_orderable(select_kth(v, qt, M, bin, width, n_edges, (unsigned)lo, s_hist, s_pick));
    const float b = from_orderable(select_kth(v, qt, M, bin, width, n_edges, (unsigned)hi, s_hist, s_pick));
    const double d = (double)b - (double)a;                    // numpy's _lerp, including its t >= 0.5 form
    return (float)(g >= 0.5 ? (double)b - d * (1.0 - g) : (double)a + d * g);
}

// out [K,5,n_bins+1]: u_perp_resolution, u_perp_scaled_resolution, u_par_resolution, u_par_scaled_resolution, R
// (the layout of np.histogram(qT_hist, bins=40, range=(0,400), weights=...)[0]: bin centres land in bins 0..38, bin 39 = 0)
__global__ void __launch_bounds__(RS_THREADS) k_resolution_summary(const float *__restrict__ vals, long long ld,
                                                                   const float *__restrict__ qt, int M, float width,
                                                                   int n_bins, float *__restrict__ out) {
    __shared__ unsigned s_hist[256], s_pick[2];
    __shared__ double s_red[RS_THREADS];
    __shared__ int s_cnt[RS_THREADS];
    const int bin = blockIdx.x + 1, key = blockIdx.y, n_edges = n_bins + 1;
    const float *u_perp = vals + ((size_t)key * 3 + 0) * ld, *u_par = vals + ((size_t)key * 3 + 1) * ld,
                *resp = vals + ((size_t)key * 3 + 2) * ld;
    double sum = 0.0;
    int cnt = 0;
    for (int i = threadIdx.x; i < M; i += RS_THREADS)
        if (digitize(qt[i], width, n_edges) == bin) { sum += (double)resp[i]; ++cnt; }
    s_red[threadIdx.x] = sum; s_cnt[threadIdx.x] = cnt;
    __syncthreads();
    for (int s = RS_THREADS / 2; s > 0; s >>= 1) {
        if (threadIdx.x < s) { s_red[threadIdx.x] += s_red[threadIdx.x + s]; s_cnt[threadIdx.x] += s_cnt[threadIdx.x + s]; }
        __syncthreads();
    }
    const int n = s_cnt[0];
    const float mean_r = n > 0 ? (float)(s_red[0] / (double)n) : __int_as_float(0x7fc00000);
    __syncthreads();
    float r[5];
    if (n > 0) {
        const float p84 = quantile(u_perp, qt, M, bin, width, n_edges, n, 0.84, s_hist, s_pick);
        const float p16 = quantile(u_perp, qt, M, bin, width, n_edges, n, 0.16, s_hist, s_pick);
        const float a84 = quantile(u_par, qt, M, bin, width, n_edges, n, 0.84, s_hist, s_pick);
        const float a16 = quantile(u_par, qt, M, bin, width, n_edges, n, 0.16, s_hist, s_pick);
        r[0] = (p84 - p16) / 2.f;
        r[2] = (a84 - a16) / 2.f;
        // quantiles of u / mean(R): dividing by a positive number keeps the order, by a negative one reverses it
        // (q84 of the scaled values is then q16 of the unscaled ones over mean(R)); either way (q84 - q16)/2 scales by 1/|mean R|
        r[1] = mean_r > 0.f ? (p84 / mean_r - p16 / mean_r) / 2.f : (p16 / mean_r - p84 / mean_r) / 2.f;
        r[3] = mean_r > 0.f ? (a84 / mean_r - a16 / mean_r) / 2.f : (a16 / mean_r - a84 / mean_r) / 2.f;
        r[4] = mean_r;
    } else {
#pragma unroll
        for (int q = 0; q < 5; ++q) r[q] = __int_as_float(0x7fc00000);
    }
    if (threadIdx.x == 0) {
        const int W = n_bins + 1;
#pragma unroll
        for (int q = 0; q < 5; ++q) {
            out[((size_t)key * 5 + q) * W + (bin - 1)] = r[q];
            if (bin == 1) out[((size_t)key * 5 + q) * W + n_bins] = 0.f;
        }
    }
}

}  // namespace
}  // namespace gm

using namespace gm;

extern "C" {

int gm_resolution_summary(const float *vals, long long ld, const float *qt, int M, int K, float bin_width, int n_bins,
                          float *out, void *stream) {
    if (M < 0 || K < 1 || n_bins < 1 || !(bin_width > 0.f) || !out || (M > 0 && (!vals || !qt || ld < M)))
        return GM_ERR_INVALID_ARGUMENT;
    dim3 grid(n_bins, K);
    k_resolution_summary<<<grid, RS_THREADS, 0, as_stream(stream)>>>(vals, ld, qt, M, bin_width, n_bins, out);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

}  // extern "C"
