// Live measurements of two on-chip peaks bench.py needs as roofline denominators and MEASURED_PEAKS.json does not carry:
// the shared-memory read bandwidth (what physically feeds k_geo_aggr: E x 128 B of staged-row reads per launch) and the
// FFMA stream rate (the k-NN search is FSUB + FFMA bound).  Both synchronise (they are measurements, not part of the path).
#include "common.cuh"

namespace gm {
namespace {

constexpr int MB_THREADS = 1024;
constexpr int MB_SMEM_F4 = 2048;          // 32 KB per CTA, 2 CTAs per SM

__global__ void __launch_bounds__(MB_THREADS, 2) k_smem_read(int iters, float *sink) {
    __shared__ float4 buf[MB_SMEM_F4];
    for (int i = threadIdx.x; i < MB_SMEM_F4; i += MB_THREADS) buf[i] = make_float4((float)i, 1.f, 2.f, 3.f);
    __syncthreads();
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    unsigned idx = threadIdx.x;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {     // consecutive lanes read consecutive 16-byte words: conflict-free LDS.128
            const float4 q = buf[(idx + u * 131) & (MB_SMEM_F4 - 1)];
            acc.x += q.x; acc.y += q.y; acc.z += q.z; acc.w += q.w;
        }
        idx += 257;
    }
    if (acc.x + acc.y + acc.z + acc.w == -1.f) sink[0] = acc.x;
}

__global__ void __launch_bounds__(MB_THREADS, 2) k_ffma(int iters, float a, float b, float *sink) {
    float acc[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) acc[k] = (float)(threadIdx.x + k);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int k = 0; k < 16; ++k) acc[k] = fmaf(acc[k], a, b);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 16; ++k) s += acc[k];
    if (s == -1.f) sink[0] = s;
}

__global__ void __launch_bounds__(MB_THREADS, 2) k_ffma2(int iters, float a, float b, float *sink) {
    float2 acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = make_float2((float)(threadIdx.x + k), (float)k);
    const float2 a2 = make_float2(a, a), b2 = make_float2(b, b);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] = ffma2(acc[k], a2, b2);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += acc[k].x + acc[k].y;
    if (s == -1.f) sink[0] = s;
}

template <typename F>
int timed(F &&launch, float *ms_out) {
    cudaEvent_t e0, e1;
    GM_CUDA_TRY(cudaEventCreate(&e0));
    GM_CUDA_TRY(cudaEventCreate(&e1));
    launch();                                  // warm-up
    GM_CUDA_TRY(cudaEventRecord(e0, 0));
    launch();
    GM_CUDA_TRY(cudaEventRecord(e1, 0));
    GM_CUDA_TRY(cudaEventSynchronize(e1));
    GM_CUDA_TRY(cudaEventElapsedTime(ms_out, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    GM_CUDA_TRY(cudaGetLastError());
    return GM_OK;
}

}  // namespace
}  // namespace gm

using namespace gm;

extern "C" {

int gm_bench_smem_read(int iters, float *ms_out, double *gbytes_out) {
    if (iters < 1 || !ms_out || !gbytes_out) return GM_ERR_INVALID_ARGUMENT;
    const int grid = num_sms() * 2;
    float *sink = nullptr;
    GM_CUDA_TRY(cudaMalloc(&sink, sizeof(float)));
    const int rc = timed([&] { k_smem_read<<<grid, MB_THREADS>>>(iters * 64, sink); }, ms_out);
    cudaFree(sink);
    *gbytes_out = (double)grid * MB_THREADS * (double)iters * 64.0 * 8.0 * 16.0 / 1e9;
    return rc;
}

int gm_bench_ffma(int iters, float *ms_out, double *gflop_out) {
    if (iters < 1 || !ms_out || !gflop_out) return GM_ERR_INVALID_ARGUMENT;
    const int grid = num_sms() * 2;
    float *sink = nullptr;
    GM_CUDA_TRY(cudaMalloc(&sink, sizeof(float)));
    const int rc = timed([&] { k_ffma<<<grid, MB_THREADS>>>(iters * 16, 0.999f, 0.001f, sink); }, ms_out);
    cudaFree(sink);
    *gflop_out = (double)grid * MB_THREADS * (double)iters * 16.0 * 128.0 * 2.0 / 1e9;
    return rc;
}

/* the packed form: fma.rn.f32x2 (FFMA2), 4 flop per instruction and lane */
int gm_bench_ffma2(int iters, float *ms_out, double *gflop_out) {
    if (iters < 1 || !ms_out || !gflop_out) return GM_ERR_INVALID_ARGUMENT;
    const int grid = num_sms() * 2;
    float *sink = nullptr;
    GM_CUDA_TRY(cudaMalloc(&sink, sizeof(float)));
    const int rc = timed([&] { k_ffma2<<<grid, MB_THREADS>>>(iters * 16, 0.999f, 0.001f, sink); }, ms_out);
    cudaFree(sink);
    *gflop_out = (double)grid * MB_THREADS * (double)iters * 16.0 * 64.0 * 4.0 / 1e9;
    return rc;
}

}  // extern "C"
