// Shared device/host helpers for libgraphmet.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "graphmet.h"

namespace gm {

constexpr int H = GM_HIDDEN;       // hidden_dim of model/net.py:41-43
constexpr int WARP = 32;

extern int g_last_cuda_error;
extern long long g_launch_count;      // kernels launched by this library (gm_launch_count)

inline int cuda_fail(cudaError_t e) {
    g_last_cuda_error = (int)e;
    return GM_ERR_CUDA;
}

#define GM_CUDA_TRY(expr)                                   \
    do {                                                    \
        cudaError_t _e = (expr);                            \
        if (_e != cudaSuccess) return ::gm::cuda_fail(_e);  \
    } while (0)

#define GM_LAUNCH_CHECK()                  \
    do {                                   \
        ++::gm::g_launch_count;            \
        GM_CUDA_TRY(cudaGetLastError());   \
    } while (0)

#define GM_TRY(expr)                 \
    do {                             \
        int _s = (expr);             \
        if (_s != GM_OK) return _s;  \
    } while (0)

int num_sms();

// Optional per-kernel-family CUDA-event timing on the launching stream (gm_timing_enable / gm_timing_read).
struct TimedRegion {
    int slot;
    cudaStream_t st;
    bool on;
    TimedRegion(int slot, cudaStream_t st);
    ~TimedRegion();
};

inline cudaStream_t as_stream(void *s) { return reinterpret_cast<cudaStream_t>(s); }

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Grid for a persistent / grid-stride kernel: enough CTAs to cover `work_items / per_cta`,
// capped at ctas_per_sm * #SM so the grid is a whole number of resident waves.
inline int persistent_grid(long long work_ctas, int ctas_per_sm) {
    long long cap = (long long)num_sms() * ctas_per_sm;
    if (work_ctas < 1) work_ctas = 1;
    return (int)(work_ctas < cap ? work_ctas : cap);
}

// Resident CTAs per SM of a kernel at a launch configuration (occupancy query; falls back to `fallback`).
template <typename K>
inline int resident_ctas(K kernel, int threads, size_t smem, int fallback) {
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, threads, smem) != cudaSuccess || n < 1) {
        cudaGetLastError();
        n = fallback;
    }
    return n;
}

// ----------------------------------------------------------------------------- device
__device__ __forceinline__ float elu(float v) { return v > 0.f ? v : expm1f(v); }
// derivative of ELU expressed through its OUTPUT a = elu(z): z>0 -> 1, else a + 1
__device__ __forceinline__ float elu_grad_from_out(float a) { return a > 0.f ? 1.f : a + 1.f; }
__device__ __forceinline__ float sigmoidf(float v) { return 1.f / (1.f + expf(-v)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int warp_sum_int(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int warp_excl_scan_int(int v, int lane, int &total) {
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    total = __shfl_sync(0xffffffffu, inc, 31);
    return inc - v;
}

// Welford/Chan moments of one channel: (n, mean, M2). Combination is done in a fixed order
// everywhere so BatchNorm statistics are bit-deterministic run to run.
struct Moments {
    float n, mean, m2;
};
__device__ __forceinline__ Moments moments_combine(Moments a, Moments b) {
    if (b.n == 0.f) return a;
    if (a.n == 0.f) return b;
    float n = a.n + b.n;
    float d = b.mean - a.mean;
    float f = b.n / n;
    Moments r;
    r.n = n;
    r.mean = a.mean + d * f;
    r.m2 = a.m2 + b.m2 + d * d * a.n * f;
    return r;
}

// y[o] = b[o] + sum_k W[o][k] x[k], W row-major [OUT][IN] in shared memory (16B aligned rows when
// IN % 4 == 0); every thread of the warp reads the same weight address => LDS broadcast.
template <int IN, int OUT>
__device__ __forceinline__ void dense_smem(const float *__restrict__ W, const float *__restrict__ b,
                                           const float (&x)[IN], float (&y)[OUT]) {
#pragma unroll
    for (int o = 0; o < OUT; ++o) {
        float acc = b[o];
        if constexpr (IN % 4 == 0) {
            const float4 *w4 = reinterpret_cast<const float4 *>(W + o * IN);
#pragma unroll
            for (int k = 0; k < IN / 4; ++k) {
                float4 w = w4[k];
                acc = fmaf(w.x, x[4 * k + 0], acc);
                acc = fmaf(w.y, x[4 * k + 1], acc);
                acc = fmaf(w.z, x[4 * k + 2], acc);
                acc = fmaf(w.w, x[4 * k + 3], acc);
            }
        } else {
#pragma unroll
            for (int k = 0; k < IN; ++k) acc = fmaf(W[o * IN + k], x[k], acc);
        }
        y[o] = acc;
        // keep ptxas from hoisting every weight load of the unrolled product to the top (register spills)
        if ((o & 3) == 3) asm volatile("" ::: "memory");
    }
}

// y[k] = sum_o W[o][k] g[o]  (transpose product, for input gradients)
template <int IN, int OUT>
__device__ __forceinline__ void dense_t_smem(const float *__restrict__ W, const float (&g)[OUT],
                                             float (&y)[IN]) {
#pragma unroll
    for (int k = 0; k < IN; ++k) y[k] = 0.f;
#pragma unroll
    for (int o = 0; o < OUT; ++o) {
        if constexpr (IN % 4 == 0) {
            const float4 *w4 = reinterpret_cast<const float4 *>(W + o * IN);
#pragma unroll
            for (int k = 0; k < IN / 4; ++k) {
                float4 w = w4[k];
                y[4 * k + 0] = fmaf(w.x, g[o], y[4 * k + 0]);
                y[4 * k + 1] = fmaf(w.y, g[o], y[4 * k + 1]);
                y[4 * k + 2] = fmaf(w.z, g[o], y[4 * k + 2]);
                y[4 * k + 3] = fmaf(w.w, g[o], y[4 * k + 3]);
            }
        } else {
#pragma unroll
            for (int k = 0; k < IN; ++k) y[k] = fmaf(W[o * IN + k], g[o], y[k]);
        }
        if ((o & 3) == 3) asm volatile("" ::: "memory");
    }
}

// reference pdg remap: model/graph_met_network.py:52-54 (sequential where-chain over
// [1,2,11,13,22,130,211]; any other value passes through unchanged)
__device__ __forceinline__ long long pdg_remap(long long pdg) {
    long long v = pdg < 0 ? -pdg : pdg;
    const int pdgs[7] = {1, 2, 11, 13, 22, 130, 211};
#pragma unroll
    for (int i = 0; i < 7; ++i) v = (v == pdgs[i]) ? i : v;
    return v;
}

}  // namespace gm
