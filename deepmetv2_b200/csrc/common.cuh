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

// Programmatic dependent launch (griddepcontrol): a kernel launched with GM_LAUNCH_PDL may begin -- block scheduling,
// barrier / TMEM setup, staging of WEIGHTS into shared memory -- while its predecessor in the stream is still running;
// it must execute pdl_wait() before it touches anything the predecessor wrote, and it lets its own successor start
// with pdl_trigger() (always AFTER its pdl_wait(), so at most two kernels overlap and a prologue may read whatever
// was complete two kernels back: parameters, never activations or statistics).  At DY/TT batch sizes a training step is
// ~27 dependent launches of 15-70 us: the launch gaps and per-CTA prologues were ~10 % of it.  GM_PDL=0 disables.
bool pdl_enabled();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args &&...args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#define GM_LAUNCH_PDL(kernel, grid, block, smem, st, ...)                                       \
    GM_CUDA_TRY(::gm::launch_pdl(kernel, dim3(grid), dim3(block), smem, st, __VA_ARGS__))   /* then GM_LAUNCH_CHECK() */

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
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_sync() { pdl_wait(); pdl_trigger(); }      // kernels without a prologue worth overlapping

// What the LAST kernel of a training step mirrors into pinned (device-mapped) host memory so the host can see device
// error flags one step late without a copy or a sync: {step, GM_FLAG_* of the forward, flags of the graph build}.
struct StepStatus {
    const int *err_flags;     // device int[1] (gm_net_forward's err_flags) or NULL
    const int *graph_stats;   // device int[4] (stats of the graph build; [3] = flags) or NULL
    int *host;                // pinned host int[4] or NULL
    int step;
    __device__ __forceinline__ void publish() const {
        if (!host) return;
        volatile int *h = host;
        h[1] = err_flags ? err_flags[0] : 0;
        h[2] = graph_stats ? graph_stats[3] : 0;
        __threadfence_system();
        h[0] = step;
    }
};

// Packed fp32x2 arithmetic (sm_100: FFMA2 / FADD2 -- one instruction, two IEEE round-to-nearest results, bit-identical
// to the scalar fmaf / subtraction on each half).
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    unsigned long long d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;"
        : "=l"(d)
        : "l"(*reinterpret_cast<unsigned long long *>(&a)), "l"(*reinterpret_cast<unsigned long long *>(&b)),
          "l"(*reinterpret_cast<unsigned long long *>(&c)));
    return *reinterpret_cast<float2 *>(&d);
}
__device__ __forceinline__ float2 fsub2(float2 a, float2 b) {
    unsigned long long d;
    asm("sub.rn.f32x2 %0, %1, %2;"
        : "=l"(d)
        : "l"(*reinterpret_cast<unsigned long long *>(&a)), "l"(*reinterpret_cast<unsigned long long *>(&b)));
    return *reinterpret_cast<float2 *>(&d);
}

__device__ __forceinline__ float elu(float v) { return v > 0.f ? v : expm1f(v); }
// ELU for the reduced-precision (bf16 / tf32 operand) paths only: expm1 as MUFU.EX2 - 1 for v <= -0.25 (absolute error
// < 3e-7, relative < 1.5e-6) and a degree-6 Taylor polynomial above (relative error < 1e-7): ~14 instructions instead
// of expm1f's ~32.  The fp32 parity paths keep expm1f.
__device__ __forceinline__ float elu_fast(float v) {
    const float t = v * (1.f + v * (0.5f + v * (0.16666667f + v * (0.041666668f + v * (0.008333334f + v * 0.0013888889f)))));
    float e;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(v * 1.4426950408889634f));
    return v > 0.f ? v : (v > -0.25f ? t : e - 1.f);
}
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

// Register-blocked tile product for the node-level layers: a CTA of 32*(O/4) threads multiplies a tile of 128 node
// rows held in shared memory ([128][LD] floats, LD odd so row-per-lane access is conflict-free) by a K x O weight
// matrix stored K-MAJOR in shared memory (wT[k*O + o], 16-byte aligned).  Thread t owns the 4 x 4 block
//     nodes  ng, ng+32, ng+64, ng+96  (ng = t / (O/4))    x    outputs 4*og .. 4*og+3  (og = t % (O/4)),
// so per k it issues 4 scalar LDS (x) + 1 LDS.128 (w) for 16 independent FMAs, and a warp's results are whole
// contiguous output rows (coalesced float4 stores straight from registers).  acc[i][j] += x[ng+32i][k] * wT[k][4og+j].
// PITCH (>= O, floats): row pitch of wT when the O outputs are a column block of a wider K-major matrix.
template <int K, int O, int LD, int PITCH = O>
__device__ __forceinline__ void tile_gemm(const float *__restrict__ sX, const float *__restrict__ wT, int ng, int og,
                                          float (&acc)[4][4]) {
    const float *x0 = sX + ng * LD;
    const float4 *w4 = reinterpret_cast<const float4 *>(wT) + og;
#pragma unroll 8
    for (int k = 0; k < K; ++k) {
        const float4 w = w4[k * (PITCH / 4)];
        const float a0 = x0[k], a1 = x0[32 * LD + k], a2 = x0[64 * LD + k], a3 = x0[96 * LD + k];
        acc[0][0] = fmaf(a0, w.x, acc[0][0]); acc[0][1] = fmaf(a0, w.y, acc[0][1]);
        acc[0][2] = fmaf(a0, w.z, acc[0][2]); acc[0][3] = fmaf(a0, w.w, acc[0][3]);
        acc[1][0] = fmaf(a1, w.x, acc[1][0]); acc[1][1] = fmaf(a1, w.y, acc[1][1]);
        acc[1][2] = fmaf(a1, w.z, acc[1][2]); acc[1][3] = fmaf(a1, w.w, acc[1][3]);
        acc[2][0] = fmaf(a2, w.x, acc[2][0]); acc[2][1] = fmaf(a2, w.y, acc[2][1]);
        acc[2][2] = fmaf(a2, w.z, acc[2][2]); acc[2][3] = fmaf(a2, w.w, acc[2][3]);
        acc[3][0] = fmaf(a3, w.x, acc[3][0]); acc[3][1] = fmaf(a3, w.y, acc[3][1]);
        acc[3][2] = fmaf(a3, w.z, acc[3][2]); acc[3][3] = fmaf(a3, w.w, acc[3][3]);
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
