// e. Event-sharded data parallelism: the gradient all-reduce FUSED with the optimiser, over NVLink peer memory.
//
// The reference trains on one device (train.py:72); what DistributedDataParallel + optimizer.step() would do around
// train.py:51-52 is here ONE kernel per step.  The payload is tiny (6,641 gradient floats + 2 scalars = 26.6 KB), so
// the collective is pure latency: a NCCL all-reduce issued from Python cost +23 us / +50 us per step at 4 / 8 GPUs
// (side stream, two event hops, the PDL chain broken before k_adamw).  Here every rank STORES its slice straight
// into a slot of every peer's exchange buffer (cudaIpc-mapped, NVLink / NVSwitch), raises a per-(rank, CTA) flag with
// release semantics, waits for the matching slices, adds them IN RANK ORDER -- every rank forms the same sum bit for
// bit, so the replicas never drift -- and applies AdamW.  One-shot all-reduce: world^2 x 26.6 KB cross the switch per
// step (1.5 MB at 8 GPUs: ~2 us of NVLink time), one NVLink round trip of latency, no extra launch.
//
//   exchange buffer of a rank:   float slots[2][world][NP]        (parity of the call, source rank, payload)
//                                unsigned flags[world][CM_MAX_CTAS] (source rank, CTA) = sequence number of the call
//   payload:                     grads[0..n) | local_events | local_events * local_loss | pad to a multiple of 4
//
// Why two slot parities are enough: rank A can start call s+1 (and overwrite slots[(s+1)&1]) while rank B still reads
// slots[s&1] of call s, but A cannot reach call s+2 before B has pushed call s+1, which B does after its call s.
#include "common.cuh"

struct GmComm {
    int rank, world, n, np;
    unsigned seq;
    size_t bytes;
    int device;
    void *local;
    void *peer[GM_COMM_MAX_WORLD];
    bool mapped[GM_COMM_MAX_WORLD];
    bool connected;
};

namespace gm {
namespace {

constexpr int CM_THREADS = 256;
constexpr int CM_MAX_CTAS = 64;                       // payload <= 16384 floats
constexpr unsigned long long CM_TIMEOUT_NS = 20ull * 1000 * 1000 * 1000;

struct CommView {
    float *slots[GM_COMM_MAX_WORLD];                  // every rank's slots[2][world][np]
    unsigned *flags[GM_COMM_MAX_WORLD];               // every rank's flags[world][CM_MAX_CTAS]
    int rank, world, np;
    unsigned seq;
};

inline size_t slots_bytes(int np, int world) { return align_up((size_t)2 * world * np * sizeof(float), 256); }
inline size_t comm_bytes(int np, int world) { return slots_bytes(np, world) + (size_t)world * CM_MAX_CTAS * sizeof(unsigned); }

__device__ __forceinline__ void st_release_sys(unsigned *p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float ld_relaxed_sys(const float *p) {
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long globaltimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// bounded wait for a flag written by a peer: a dead peer must trap this kernel, not hang the GPU
__device__ __forceinline__ void wait_flag(const unsigned *p, unsigned seq) {
    if ((int)(ld_acquire_sys(p) - seq) >= 0) return;
    const unsigned long long t0 = globaltimer();
    for (;;) {
        __nanosleep(40);
        if ((int)(ld_acquire_sys(p) - seq) >= 0) return;
        if (globaltimer() - t0 > CM_TIMEOUT_NS) __trap();
    }
}

// One element of the payload per thread; CTA b owns elements [256 b, 256 b + 256).  The CTA that owns element n
// (the event count) is the "tail" CTA: every CTA also waits for the tail slice of every rank, because the AdamW
// update of any element needs the global event count.
__global__ void __launch_bounds__(CM_THREADS)
k_exchange_adamw(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m, float *__restrict__ v, int n,
                 float lr, float b1, float b2, float eps, float wd, float bc1, float bc2_sqrt, float local_events,
                 const float *__restrict__ local_loss, CommView c, float *__restrict__ global_loss, StepStatus status) {
    pdl_wait();                                          // g, local_loss: the previous kernels' outputs; no early trigger
    const int t = threadIdx.x, i = blockIdx.x * CM_THREADS + t;
    const int W = c.world, par = (int)(c.seq & 1u);
    const int tail = n / CM_THREADS;
    if (blockIdx.x == 0 && t == 0) status.publish();
    // ---- push my slice into slot `rank` of every rank's buffer (my own included: one code path, one summation order)
    if (i < c.np) {
        float mine = 0.f;
        if (i < n) mine = g[i];
        else if (i == n) mine = local_events;
        else if (i == n + 1) mine = local_events > 0.f ? local_events * local_loss[0] : 0.f;
        const size_t off = ((size_t)par * W + c.rank) * c.np + i;
        for (int r = 0; r < W; ++r) c.slots[(c.rank + r) % W][off] = mine;     // start with myself, then round-robin peers
    }
    __threadfence_system();
    __syncthreads();
    if (t < W) st_release_sys(c.flags[t] + (size_t)c.rank * CM_MAX_CTAS + blockIdx.x, c.seq);
    // ---- wait for the same slice (and the tail slice) of every rank
    const unsigned *my_flags = c.flags[c.rank];
    if (t < W) wait_flag(my_flags + (size_t)t * CM_MAX_CTAS + blockIdx.x, c.seq);
    else if (t >= 32 && t < 32 + W && (int)blockIdx.x != tail) wait_flag(my_flags + (size_t)(t - 32) * CM_MAX_CTAS + tail, c.seq);
    __syncthreads();
    const float *mine_slots = c.slots[c.rank] + (size_t)par * W * c.np;
    float events = 0.f;
    for (int r = 0; r < W; ++r) events += ld_relaxed_sys(mine_slots + (size_t)r * c.np + n);
    const float inv_events = 1.f / fmaxf(events, 1.f);
    if (i < n) {
        float gi = 0.f;
        for (int r = 0; r < W; ++r) gi += ld_relaxed_sys(mine_slots + (size_t)r * c.np + i);
        gi *= inv_events;
        float pi = p[i] * (1.f - lr * wd);
        const float mi = b1 * m[i] + (1.f - b1) * gi;
        const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
        const float denom = sqrtf(vi) / bc2_sqrt + eps;
        pi -= (lr / bc1) * (mi / denom);
        p[i] = pi; m[i] = mi; v[i] = vi;
    } else if (i == n + 1 && global_loss) {
        float ls = 0.f;
        for (int r = 0; r < W; ++r) ls += ld_relaxed_sys(mine_slots + (size_t)r * c.np + i);
        global_loss[0] = ls * inv_events;
    }
}

// The same exchange as TWO kernels, so that the wait for the peers can be moved behind work of the NEXT step that does
// not depend on the updated weights (its radius-graph build): k_exchange_post ends step s (push + flags, never waits),
// k_exchange_apply runs in step s + 1 right before the forward (wait, rank-ordered sum, AdamW).  Same slots, flags,
// parity and summation order as the fused kernel: the parameters are bit-identical either way.
__global__ void __launch_bounds__(CM_THREADS)
k_exchange_post(const float *__restrict__ g, int n, float local_events, const float *__restrict__ local_loss, CommView c) {
    pdl_wait();
    const int t = threadIdx.x, i = blockIdx.x * CM_THREADS + t;
    const int W = c.world, par = (int)(c.seq & 1u);
    if (i < c.np) {
        float mine = 0.f;
        if (i < n) mine = g[i];
        else if (i == n) mine = local_events;
        else if (i == n + 1) mine = local_events > 0.f ? local_events * local_loss[0] : 0.f;
        const size_t off = ((size_t)par * W + c.rank) * c.np + i;
        for (int r = 0; r < W; ++r) c.slots[(c.rank + r) % W][off] = mine;
    }
    __threadfence_system();
    __syncthreads();
    if (t < W) st_release_sys(c.flags[t] + (size_t)c.rank * CM_MAX_CTAS + blockIdx.x, c.seq);
}

__global__ void __launch_bounds__(CM_THREADS)
k_exchange_apply(float *__restrict__ p, float *__restrict__ m, float *__restrict__ v, int n, float lr, float b1, float b2,
                 float eps, float wd, float bc1, float bc2_sqrt, CommView c, float *__restrict__ global_loss, StepStatus status) {
    pdl_wait();
    const int t = threadIdx.x, i = blockIdx.x * CM_THREADS + t;
    const int W = c.world, par = (int)(c.seq & 1u);
    const int tail = n / CM_THREADS;
    if (blockIdx.x == 0 && t == 0) status.publish();
    const unsigned *my_flags = c.flags[c.rank];
    if (t < W) wait_flag(my_flags + (size_t)t * CM_MAX_CTAS + blockIdx.x, c.seq);
    else if (t >= 32 && t < 32 + W && (int)blockIdx.x != tail) wait_flag(my_flags + (size_t)(t - 32) * CM_MAX_CTAS + tail, c.seq);
    __syncthreads();
    const float *mine_slots = c.slots[c.rank] + (size_t)par * W * c.np;
    float events = 0.f;
    for (int r = 0; r < W; ++r) events += ld_relaxed_sys(mine_slots + (size_t)r * c.np + n);
    const float inv_events = 1.f / fmaxf(events, 1.f);
    if (i < n) {
        float gi = 0.f;
        for (int r = 0; r < W; ++r) gi += ld_relaxed_sys(mine_slots + (size_t)r * c.np + i);
        gi *= inv_events;
        float pi = p[i] * (1.f - lr * wd);
        const float mi = b1 * m[i] + (1.f - b1) * gi;
        const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
        const float denom = sqrtf(vi) / bc2_sqrt + eps;
        pi -= (lr / bc1) * (mi / denom);
        p[i] = pi; m[i] = mi; v[i] = vi;
    } else if (i == n + 1 && global_loss) {
        float ls = 0.f;
        for (int r = 0; r < W; ++r) ls += ld_relaxed_sys(mine_slots + (size_t)r * c.np + i);
        global_loss[0] = ls * inv_events;
    }
}

// comm == NULL: the same arithmetic without peers
__global__ void __launch_bounds__(CM_THREADS)
k_local_adamw(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m, float *__restrict__ v, int n,
              float lr, float b1, float b2, float eps, float wd, float bc1, float bc2_sqrt, float local_events,
              const float *__restrict__ local_loss, float *__restrict__ global_loss, StepStatus status) {
    pdl_wait();
    const int i = blockIdx.x * CM_THREADS + threadIdx.x;
    if (i == 0) {
        status.publish();
        if (global_loss) global_loss[0] = local_events > 0.f ? local_loss[0] : 0.f;
    }
    const float inv_events = 1.f / fmaxf(local_events, 1.f);
    if (i < n) {
        const float gi = g[i] * inv_events;
        float pi = p[i] * (1.f - lr * wd);
        const float mi = b1 * m[i] + (1.f - b1) * gi;
        const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
        const float denom = sqrtf(vi) / bc2_sqrt + eps;
        pi -= (lr / bc1) * (mi / denom);
        p[i] = pi; m[i] = mi; v[i] = vi;
    }
}

}  // namespace
}  // namespace gm

using namespace gm;

extern "C" {

int gm_comm_create(int n_floats, int rank, int world, GmComm **comm_out, void *ipc_handle_out) {
    if (!comm_out || !ipc_handle_out || n_floats < 1 || world < 1 || rank < 0 || rank >= world) return GM_ERR_INVALID_ARGUMENT;
    static_assert(sizeof(cudaIpcMemHandle_t) == GM_COMM_HANDLE_BYTES, "IPC handle size");
    const int np = (n_floats + 2 + 3) & ~3;
    if (world > GM_COMM_MAX_WORLD || (np + CM_THREADS - 1) / CM_THREADS > CM_MAX_CTAS) return GM_ERR_UNSUPPORTED;
    GmComm *c = new GmComm();
    c->rank = rank; c->world = world; c->n = n_floats; c->np = np; c->seq = 0; c->connected = false;
    c->bytes = comm_bytes(np, world);
    for (int r = 0; r < GM_COMM_MAX_WORLD; ++r) { c->peer[r] = nullptr; c->mapped[r] = false; }
    cudaError_t e = cudaGetDevice(&c->device);
    if (e == cudaSuccess) e = cudaMalloc(&c->local, c->bytes);
    if (e == cudaSuccess) e = cudaMemset(c->local, 0, c->bytes);
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t *>(ipc_handle_out), c->local);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        if (c->local) cudaFree(c->local);
        delete c;
        cudaGetLastError();
        return cuda_fail(e);
    }
    c->peer[rank] = c->local;
    *comm_out = c;
    return GM_OK;
}

int gm_comm_connect(GmComm *c, const void *all_handles) {
    if (!c || !all_handles) return GM_ERR_INVALID_ARGUMENT;
    const cudaIpcMemHandle_t *h = reinterpret_cast<const cudaIpcMemHandle_t *>(all_handles);
    for (int r = 0; r < c->world; ++r) {
        if (r == c->rank || c->mapped[r]) continue;
        void *ptr = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&ptr, h[r], cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) { cudaGetLastError(); return cuda_fail(e); }
        c->peer[r] = ptr;
        c->mapped[r] = true;
    }
    c->connected = true;
    return GM_OK;
}

int gm_comm_destroy(GmComm *c) {
    if (!c) return GM_OK;
    cudaDeviceSynchronize();
    for (int r = 0; r < c->world; ++r)
        if (c->mapped[r]) cudaIpcCloseMemHandle(c->peer[r]);
    if (c->local) cudaFree(c->local);
    cudaGetLastError();
    delete c;
    return GM_OK;
}

int gm_allreduce_adamw_step(float *params, const float *grads, float *exp_avg, float *exp_avg_sq, int n, float lr,
                            float beta1, float beta2, float eps, float weight_decay, int step, float local_events,
                            const float *local_loss_dev, GmComm *comm, float *global_loss_out, const int *err_flags_dev,
                            const int *graph_stats_dev, int *status_host, void *stream) {
    if (n < 1 || step < 1 || !params || !grads || !exp_avg || !exp_avg_sq || !local_loss_dev || local_events < 0.f)
        return GM_ERR_INVALID_ARGUMENT;
    const double bc1 = 1.0 - pow((double)beta1, (double)step);
    const double bc2 = 1.0 - pow((double)beta2, (double)step);
    const StepStatus status{err_flags_dev, graph_stats_dev, status_host, step};
    cudaStream_t st = as_stream(stream);
    if (!comm || comm->world == 1) {
        GM_LAUNCH_PDL(k_local_adamw, (n + CM_THREADS - 1) / CM_THREADS, CM_THREADS, 0, st, params, grads, exp_avg, exp_avg_sq, n,
                      lr, beta1, beta2, eps, weight_decay, (float)bc1, (float)sqrt(bc2), local_events, local_loss_dev,
                      global_loss_out, status);
        GM_LAUNCH_CHECK();
        return GM_OK;
    }
    if (!comm->connected || comm->n != n) return GM_ERR_INVALID_ARGUMENT;
    CommView v{};
    v.rank = comm->rank; v.world = comm->world; v.np = comm->np;
    v.seq = ++comm->seq;
    for (int r = 0; r < comm->world; ++r) {
        v.slots[r] = reinterpret_cast<float *>(comm->peer[r]);
        v.flags[r] = reinterpret_cast<unsigned *>(reinterpret_cast<char *>(comm->peer[r]) + slots_bytes(comm->np, comm->world));
    }
    GM_LAUNCH_PDL(k_exchange_adamw, (comm->np + CM_THREADS - 1) / CM_THREADS, CM_THREADS, 0, st, params, grads, exp_avg,
                  exp_avg_sq, n, lr, beta1, beta2, eps, weight_decay, (float)bc1, (float)sqrt(bc2), local_events,
                  local_loss_dev, v, global_loss_out, status);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

static CommView view_of(GmComm *comm, unsigned seq) {
    CommView v{};
    v.rank = comm->rank; v.world = comm->world; v.np = comm->np;
    v.seq = seq;
    for (int r = 0; r < comm->world; ++r) {
        v.slots[r] = reinterpret_cast<float *>(comm->peer[r]);
        v.flags[r] = reinterpret_cast<unsigned *>(reinterpret_cast<char *>(comm->peer[r]) + slots_bytes(comm->np, comm->world));
    }
    return v;
}

int gm_exchange_post(const float *grads, int n, float local_events, const float *local_loss_dev, GmComm *comm, void *stream) {
    if (n < 1 || !grads || !local_loss_dev || local_events < 0.f || !comm || comm->world < 2) return GM_ERR_INVALID_ARGUMENT;
    if (!comm->connected || comm->n != n) return GM_ERR_INVALID_ARGUMENT;
    const CommView v = view_of(comm, ++comm->seq);
    GM_LAUNCH_PDL(k_exchange_post, (comm->np + CM_THREADS - 1) / CM_THREADS, CM_THREADS, 0, as_stream(stream), grads, n,
                  local_events, local_loss_dev, v);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

int gm_exchange_apply_adamw(float *params, float *exp_avg, float *exp_avg_sq, int n, float lr, float beta1, float beta2,
                            float eps, float weight_decay, int step, GmComm *comm, float *global_loss_out,
                            const int *err_flags_dev, const int *graph_stats_dev, int *status_host, void *stream) {
    if (n < 1 || step < 1 || !params || !exp_avg || !exp_avg_sq || !comm || comm->world < 2) return GM_ERR_INVALID_ARGUMENT;
    if (!comm->connected || comm->n != n || comm->seq == 0) return GM_ERR_INVALID_ARGUMENT;
    const double bc1 = 1.0 - pow((double)beta1, (double)step);
    const double bc2 = 1.0 - pow((double)beta2, (double)step);
    const StepStatus status{err_flags_dev, graph_stats_dev, status_host, step};
    const CommView v = view_of(comm, comm->seq);          // the exchange the last gm_exchange_post opened
    GM_LAUNCH_PDL(k_exchange_apply, (comm->np + CM_THREADS - 1) / CM_THREADS, CM_THREADS, 0, as_stream(stream), params, exp_avg,
                  exp_avg_sq, n, lr, beta1, beta2, eps, weight_decay, (float)bc1, (float)sqrt(bc2), v, global_loss_out, status);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

}  // extern "C"
