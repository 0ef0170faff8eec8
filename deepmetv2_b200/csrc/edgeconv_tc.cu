// a5 on the 5th-generation tensor cores: EdgeConv as a per-edge dense contraction on tcgen05 with TMEM
// accumulators (sm_100a).  Replaces the gather / sub / cat / SGEMM / scatter-max sequence PyG's EdgeConv
// runs for model/graph_met_network.py:65.
//
//   out_i[c] = aggr_{j in N(i)} ( W1[c,:] x_i + W2[c,:] (x_j - x_i) + b[c] )      W = [W1 | W2]  (32 x 64)
//
// Per tile of <= 256 CSR edges (whole target rows only, so no segment ever straddles a tile or a warp):
//   producers (4 warps)  gather x_j rows (16-byte chunks, 8-16 loads in flight per thread), subtract x_i in
//                        fp32, round to tf32 / bf16 and store the K-major operand tile into shared memory;
//   MMA (1 thread)       D[m][e] = sum_k W2r[m][k] * d_e[k] with M = 128 (W2 replicated 4x so that each of the
//                        four epilogue warps owns a full copy of the 32 channels in its own TMEM lane
//                        quarter), N = edges of the tile (<= 256), K = 32; accumulator in TMEM;
//   epilogue (4 warps)   lane = output channel: walks its edge columns with tcgen05.ld and keeps a running
//                        max / arg-max / sum per target in registers -- a sequential, fixed-order,
//                        atomics-free segmented reduction -- adds p_i = W1 x_i + b and writes the row.
// Two smem stages and two 256-column TMEM accumulators overlap gather(t+1) with MMA+epilogue(t).
// Every mbarrier wait is bounded: on timeout the kernel raises a flag and traps instead of hanging.
#include "net_layout.cuh"

namespace gm {
namespace {

constexpr int TC_LMAX = 256;                 // edge rows per tile (MMA N)
constexpr int TC_TMAX = 64;                  // targets per tile
constexpr int TC_STAGES = 2;
constexpr int TC_LBO = 144;                  // bytes between the 16-byte K chunks of an 8-row core-matrix group
constexpr int TC_PRODUCERS = 128;
constexpr int TC_THREADS = 288;              // warps 0-3 epilogue, 4-7 producers, 8 MMA issuer / TMEM owner
constexpr int TC_TMEM_COLS = 512;
constexpr unsigned TC_SPIN_LIMIT = 1u << 22;

template <int PREC> struct TcCfg;
template <> struct TcCfg<GM_PREC_TF32> { static constexpr int CHUNKS = 8, KSTEPS = 4, FMT = 2; };   // 4 tf32 per chunk
template <> struct TcCfg<GM_PREC_BF16> { static constexpr int CHUNKS = 4, KSTEPS = 2, FMT = 1; };   // 8 bf16 per chunk

struct Book {                                // per-stage tile bookkeeping (written by producers)
    int i0, T, L, e0;
    int tq[5];                               // epilogue warp q owns targets [tq[q], tq[q+1])
    int pad[3];
    int rowoff[TC_TMAX + 4];                 // edge offset of target k relative to e0 (T+1 used; padded to 16 B)
    int src[TC_LMAX];                        // source node of each edge row
    unsigned lastmask[TC_LMAX / 32];         // bit set on the last edge column of every non-empty target
    unsigned char tgt[TC_LMAX];              // tile-local target of each edge row
    float p[TC_TMAX][H];                     // W1 x_i + b per target (written/read by the owning epilogue warp)
};

template <int PREC>
struct TcSmem {
    static constexpr int SBO = TcCfg<PREC>::CHUNKS * TC_LBO;
    static constexpr int A_BYTES = (128 / 8) * SBO;
    static constexpr int B_BYTES = (TC_LMAX / 8) * SBO;
    static constexpr size_t OFF_A = 0;
    static constexpr size_t OFF_B = A_BYTES;
    static constexpr size_t OFF_W1T = OFF_B + (size_t)TC_STAGES * B_BYTES;
    static constexpr size_t OFF_BIAS = OFF_W1T + H * TILE_LD * 4;
    static constexpr size_t OFF_BOOK = (OFF_BIAS + H * 4 + 15) / 16 * 16;
    static constexpr size_t OFF_RED = OFF_BOOK + TC_STAGES * sizeof(Book);
    static constexpr size_t OFF_BAR = (OFF_RED + 4 * 3 * H * 4 + 7) / 8 * 8;
    static constexpr size_t OFF_TMEM = OFF_BAR + 3 * TC_STAGES * 8;
    static constexpr size_t TOTAL = OFF_TMEM + 16;
};

// ------------------------------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// bounded wait: a protocol bug must never hang the GPU
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, int *err) {
#pragma unroll 1
    for (unsigned it = 0; it < TC_SPIN_LIMIT; ++it)
        if (mbar_try_wait(bar, parity)) return;
    if (err) atomicOr(err, 0x100);
    __trap();
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
template <int PREC>
__device__ __forceinline__ void umma(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    if constexpr (PREC == GM_PREC_TF32) {
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
            ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
    } else {
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
            ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
    }
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// K-major, no-swizzle ("interleaved") operand layout: 8-row x 16-byte core matrices; chunk stride LBO, 8-row group
// stride SBO.  Producers write 16-byte chunks; LBO = 144 makes the 8 chunks of a row hit 8 distinct bank groups.
template <int PREC>
__device__ __forceinline__ uint32_t operand_offset(int row, int chunk) {
    return (uint32_t)((row >> 3) * TcSmem<PREC>::SBO + chunk * TC_LBO + (row & 7) * 16);
}
template <int PREC>
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr) {
    uint64_t d = (uint64_t)((addr >> 4) & 0x3FFF);
    d |= (uint64_t)(TC_LBO >> 4) << 16;
    d |= (uint64_t)(TcSmem<PREC>::SBO >> 4) << 32;
    d |= 1ull << 46;                          // descriptor version (sm_100)
    return d;                                 // layout_type = 0 (SWIZZLE_NONE), base_offset = 0
}
template <int PREC>
__device__ __forceinline__ uint32_t instr_desc(int n) {
    return (1u << 4) | ((uint32_t)TcCfg<PREC>::FMT << 7) | ((uint32_t)TcCfg<PREC>::FMT << 10) |
           ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);   // F32 accum, K-major A and B, M = 128
}

__device__ __forceinline__ uint32_t to_tf32(float v) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
    return r;
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {   // lo -> lower address
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// 16 bytes of operand row from 4 (tf32) or 8 (bf16) fp32 differences
template <int PREC>
__device__ __forceinline__ void store_chunk(uint32_t addr, const float4 *xj, const float4 *xi) {
    if constexpr (PREC == GM_PREC_TF32) {
        st_shared_v4(addr, to_tf32(xj[0].x - xi[0].x), to_tf32(xj[0].y - xi[0].y), to_tf32(xj[0].z - xi[0].z),
                     to_tf32(xj[0].w - xi[0].w));
    } else {
        st_shared_v4(addr, pack_bf16(xj[0].x - xi[0].x, xj[0].y - xi[0].y), pack_bf16(xj[0].z - xi[0].z, xj[0].w - xi[0].w),
                     pack_bf16(xj[1].x - xi[1].x, xj[1].y - xi[1].y), pack_bf16(xj[1].z - xi[1].z, xj[1].w - xi[1].w));
    }
}

// first index i in [lo, hi] with a[i] >= v (a non-decreasing)
__device__ __forceinline__ int lower_bound_dev(const int *__restrict__ a, int lo, int hi, long long v) {
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if ((long long)a[mid] < v) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// ------------------------------------------------------------------------------------------ kernel
template <int PREC, int AGGR, bool ARG>
__global__ void __launch_bounds__(TC_THREADS, 1)
k_edgeconv_tc(const float *__restrict__ x, const int *__restrict__ rowptr, const int *__restrict__ col, int N,
              const float *__restrict__ W, const float *__restrict__ bias, float *__restrict__ out,
              int *__restrict__ arg, float *__restrict__ partials, int *__restrict__ err) {
    using S = TcSmem<PREC>;
    using C = TcCfg<PREC>;
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    float *s_w1t = reinterpret_cast<float *>(smem + S::OFF_W1T);
    float *s_bias = reinterpret_cast<float *>(smem + S::OFF_BIAS);
    Book *book = reinterpret_cast<Book *>(smem + S::OFF_BOOK);
    float(*s_red)[3][H] = reinterpret_cast<float(*)[3][H]>(smem + S::OFF_RED);
    uint32_t *s_tmem = reinterpret_cast<uint32_t *>(smem + S::OFF_TMEM);
    const uint32_t a_base = smem_u32(smem + S::OFF_A);
    const uint32_t b_base = smem_u32(smem + S::OFF_B);
    const uint32_t bar0 = smem_u32(smem + S::OFF_BAR);
    auto bar_full = [&](int s) { return bar0 + 8u * s; };
    auto bar_acc = [&](int s) { return bar0 + 8u * (TC_STAGES + s); };
    auto bar_empty = [&](int s) { return bar0 + 8u * (2 * TC_STAGES + s); };

    // ---- one-time setup
    if (tid == 0) {
        for (int s = 0; s < TC_STAGES; ++s) {
            mbar_init(bar_full(s), TC_PRODUCERS);
            mbar_init(bar_acc(s), 1);
            mbar_init(bar_empty(s), 4);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 8) tmem_alloc(smem_u32(s_tmem), TC_TMEM_COLS);
    // A operand: W2 (columns 32..63 of W) replicated 4x along M, rounded to the MMA input type
    for (int t = tid; t < 128 * C::CHUNKS; t += TC_THREADS) {
        const int m = t / C::CHUNKS, c = t - m * C::CHUNKS, ch = m & 31;
        const float *wrow = W + ch * CONV_IN + H;
        const uint32_t addr = a_base + operand_offset<PREC>(m, c);
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        if constexpr (PREC == GM_PREC_TF32) {
            const float4 w0 = *reinterpret_cast<const float4 *>(wrow + 4 * c);
            store_chunk<PREC>(addr, &w0, &z);
        } else {
            const float4 w01[2] = {*reinterpret_cast<const float4 *>(wrow + 8 * c),
                                   *reinterpret_cast<const float4 *>(wrow + 8 * c + 4)};
            const float4 zz[2] = {z, z};
            store_chunk<PREC>(addr, w01, zz);
        }
    }
    for (int t = tid; t < H * H; t += TC_THREADS) {
        const int c = t >> 5, k = t & 31;
        s_w1t[k * TILE_LD + c] = W[c * CONV_IN + k];
    }
    if (tid < H) s_bias[tid] = bias[tid];
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *s_tmem;

    // ---- this CTA's contiguous target range, balanced by edge count
    const long long E = rowptr[N];
    const int G = gridDim.x, b = blockIdx.x;
    const int I_beg = b == 0 ? 0 : lower_bound_dev(rowptr, 0, N, E * b / G);
    const int I_end = b == G - 1 ? N : lower_bound_dev(rowptr, 0, N, E * (b + 1) / G);

    if (warp >= 4 && warp < 8) {
        // =========================================================================== producers
        const int ptid = tid - 128;
        int i0 = I_beg;
        for (int it = 0;; ++it) {
            const int s = it & 1;
            Book &bk = book[s];
            mbar_wait(bar_empty(s), ((it >> 1) & 1) ^ 1, err);
            // -- plan the tile: whole targets, <= TC_LMAX edges, <= TC_TMAX targets
            if (warp == 4) {
                int T = 0, L = -1, e0 = 0;
                if (i0 < I_end) {
                    e0 = rowptr[i0];
                    const int ia = i0 + 1 + lane, ib = ia + 32;
                    const int va = ia <= I_end ? rowptr[ia] - e0 : 0x7fffffff;
                    const int vb = ib <= I_end ? rowptr[ib] - e0 : 0x7fffffff;
                    const unsigned ma = __ballot_sync(0xffffffffu, va <= TC_LMAX);
                    const unsigned mb = __ballot_sync(0xffffffffu, vb <= TC_LMAX);
                    T = (ma == 0xffffffffu) ? 32 + __popc(mb) : __popc(ma);   // row offsets are non-decreasing
                    if (lane < T) bk.rowoff[1 + lane] = va;
                    if (32 + lane < T) bk.rowoff[33 + lane] = vb;
                    if (lane == 0) bk.rowoff[0] = 0;
                    const int last = (T <= 32) ? __shfl_sync(0xffffffffu, va, (T - 1) & 31)
                                               : __shfl_sync(0xffffffffu, vb, (T - 33) & 31);
                    L = T > 0 ? last : -2;    // T == 0 cannot happen (degree <= 256), kept as a hard error
                }
                if (lane < TC_LMAX / 32) bk.lastmask[lane] = 0u;
                if (lane == 0) { bk.i0 = i0; bk.T = T; bk.L = L; bk.e0 = e0; bk.tq[0] = 0; bk.tq[4] = T; }
            }
            asm volatile("bar.sync 1, %0;" ::"n"(TC_PRODUCERS) : "memory");
            const int T = bk.T, L = bk.L, e0 = bk.e0;
            if (L < 0) {                       // range exhausted: publish the sentinel and leave
                mbar_arrive(bar_full(s));
                if (L == -2 && err && ptid == 0) atomicOr(err, 0x200);
                break;
            }
            // -- per-row bookkeeping: source node, tile-local target, last-column mask, warp split points
            for (int r = ptid; r < L; r += TC_PRODUCERS) {
                int lo = 0, hi = T;            // largest k with rowoff[k] <= r
                while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (bk.rowoff[mid] <= r) lo = mid; else hi = mid; }
                bk.tgt[r] = (unsigned char)lo;
                bk.src[r] = col[e0 + r];
                if (r + 1 == bk.rowoff[lo + 1]) atomicOr(&bk.lastmask[r >> 5], 1u << (r & 31));
            }
            if (ptid >= 1 && ptid <= 3) {      // first target whose row starts at or after L*q/4
                const int want = (L * ptid) >> 2;
                int lo = 0, hi = T;
                while (lo < hi) { const int mid = (lo + hi) >> 1; if (bk.rowoff[mid] < want) lo = mid + 1; else hi = mid; }
                bk.tq[ptid] = lo;
            }
            asm volatile("bar.sync 1, %0;" ::"n"(TC_PRODUCERS) : "memory");
            // -- gather x_j, subtract x_i (fp32), round, store the K-major operand rows
            const int nrows = (L + 15) & ~15;  // MMA N
            const uint32_t bs = b_base + (uint32_t)s * S::B_BYTES;
            constexpr int TPR = C::CHUNKS;                     // threads per row
            constexpr int RPI = TC_PRODUCERS / TPR;            // rows per iteration
            constexpr int F4 = 8 / TPR;                        // float4 loads per chunk (1 tf32, 2 bf16)
            const int c = ptid % TPR, rb = ptid / TPR;
#pragma unroll 1
            for (int r0 = 0; r0 < nrows; r0 += RPI * 8) {
                float4 xj[8][F4], xi[8][F4];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int r = r0 + u * RPI + rb;
                    if (r < L) {
                        const float4 *pj = reinterpret_cast<const float4 *>(x + (size_t)bk.src[r] * H) + c * F4;
#pragma unroll
                        for (int f = 0; f < F4; ++f) xj[u][f] = __ldg(pj + f);
                    } else {
#pragma unroll
                        for (int f = 0; f < F4; ++f) xj[u][f] = make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                }
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int r = r0 + u * RPI + rb;
                    if (r < L) {
                        const float4 *pi = reinterpret_cast<const float4 *>(x + (size_t)(bk.i0 + bk.tgt[r]) * H) + c * F4;
#pragma unroll
                        for (int f = 0; f < F4; ++f) xi[u][f] = __ldg(pi + f);
                    } else {
#pragma unroll
                        for (int f = 0; f < F4; ++f) xi[u][f] = make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                }
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int r = r0 + u * RPI + rb;
                    if (r < nrows) store_chunk<PREC>(bs + operand_offset<PREC>(r, c), xj[u], xi[u]);
                }
            }
            fence_async_smem();                // generic-proxy smem writes -> visible to the tensor-core (async) proxy
            mbar_arrive(bar_full(s));
            i0 += T;
        }
    } else if (warp == 8) {
        // =========================================================================== MMA issuer
        for (int it = 0;; ++it) {
            const int s = it & 1;
            mbar_wait(bar_full(s), (it >> 1) & 1, err);
            tc_fence_after();
            const int L = book[s].L;
            if (L < 0) {
                if (lane == 0) mbar_arrive(bar_acc(s));
                break;
            }
            if (L == 0) {                      // only empty rows in this tile: nothing to multiply
                if (lane == 0) mbar_arrive(bar_acc(s));
                continue;
            }
            if (lane == 0) {
                const int n = (L + 15) & ~15;
                const uint32_t idesc = instr_desc<PREC>(n);
                const uint32_t d = tmem_base + (uint32_t)(s * 256);
                const uint32_t bs = b_base + (uint32_t)s * S::B_BYTES;
#pragma unroll
                for (int ks = 0; ks < C::KSTEPS; ++ks) {
                    const uint64_t ad = smem_desc<PREC>(a_base + ks * 2 * TC_LBO);
                    const uint64_t bd = smem_desc<PREC>(bs + ks * 2 * TC_LBO);
                    umma<PREC>(d, ad, bd, idesc, ks > 0 ? 1u : 0u);
                }
                umma_commit(bar_acc(s));       // arrives when the MMAs above have completed
            }
            __syncwarp();
        }
    } else {
        // =========================================================================== epilogue (warps 0-3)
        const int q = warp;                    // TMEM lane quarter 32q..32q+31 <-> channel = lane
        Moments run{0.f, 0.f, 0.f};
        for (int it = 0;; ++it) {
            const int s = it & 1;
            Book &bk = book[s];
            mbar_wait(bar_full(s), (it >> 1) & 1, err);
            const int L = bk.L;
            if (L < 0) break;
            const int i0 = bk.i0, k_beg = bk.tq[q], k_end = bk.tq[q + 1];
            // p_k = W1 x_i + b for this warp's targets (overlaps with the MMA); empty rows are written here
            for (int k = k_beg; k < k_end; ++k) {
                const int i = i0 + k;
                const float xi = x[(size_t)i * H + lane];
                float p = s_bias[lane];
#pragma unroll
                for (int kk = 0; kk < H; ++kk) p = fmaf(s_w1t[kk * TILE_LD + lane], __shfl_sync(0xffffffffu, xi, kk), p);
                bk.p[k][lane] = p;
                if (bk.rowoff[k + 1] == bk.rowoff[k]) {          // PyG: empty rows aggregate to 0
                    out[(size_t)i * H + lane] = 0.f;
                    if (ARG) arg[(size_t)i * H + lane] = -1;
                    run.n += 1.f;
                    const float dl = 0.f - run.mean;
                    run.mean += dl / run.n;
                    run.m2 = fmaf(dl, 0.f - run.mean, run.m2);
                }
            }
            __syncwarp();
            mbar_wait(bar_acc(s), (it >> 1) & 1, err);
            tc_fence_after();
            const int c0 = bk.rowoff[k_beg], c1 = bk.rowoff[k_end];
            int k = k_beg;
            while (k < k_end && bk.rowoff[k + 1] == bk.rowoff[k]) ++k;
            float best = -INFINITY, sum = 0.f;
            int bcol = 0;
            const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(s * 256);
            for (int cb = c0 & ~31; cb < c1; cb += 32) {
                uint32_t r[32];
                tmem_ld32(tbase + (uint32_t)cb, r);
                const unsigned lm = bk.lastmask[cb >> 5];
#pragma unroll
                for (int u = 0; u < 32; ++u) {
                    const int cc = cb + u;
                    if (cc >= c0 && cc < c1) {                   // warp-uniform
                        const float v = __uint_as_float(r[u]);
                        if (AGGR == GM_AGGR_MAX) { if (v > best) { best = v; bcol = cc; } }
                        else sum += v;
                        if ((lm >> u) & 1u) {                    // last edge of target k: emit the row
                            const int i = i0 + k;
                            const float p = bk.p[k][lane];
                            const int deg = bk.rowoff[k + 1] - bk.rowoff[k];
                            float o;
                            if (AGGR == GM_AGGR_MAX) o = p + best;
                            else if (AGGR == GM_AGGR_ADD) o = fmaf((float)deg, p, sum);
                            else o = p + sum / (float)deg;
                            out[(size_t)i * H + lane] = o;
                            if (ARG) arg[(size_t)i * H + lane] = bk.src[bcol];
                            run.n += 1.f;
                            const float dl = o - run.mean;
                            run.mean += dl / run.n;
                            run.m2 = fmaf(dl, o - run.mean, run.m2);
                            best = -INFINITY; sum = 0.f; bcol = cc;
                            ++k;
                            while (k < k_end && bk.rowoff[k + 1] == bk.rowoff[k]) ++k;
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_empty(s));
        }
        s_red[q][0][lane] = run.n; s_red[q][1][lane] = run.mean; s_red[q][2][lane] = run.m2;
    }

    // ---- teardown
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 8) tmem_dealloc(tmem_base, TC_TMEM_COLS);
    if (partials && warp == 0) {
        Moments tot{0.f, 0.f, 0.f};
#pragma unroll
        for (int k = 0; k < 4; ++k) tot = moments_combine(tot, Moments{s_red[k][0][lane], s_red[k][1][lane], s_red[k][2][lane]});
        float *dst = partials + (size_t)blockIdx.x * MOM_STRIDE;
        dst[lane] = tot.n; dst[H + lane] = tot.mean; dst[2 * H + lane] = tot.m2;
    }
}

template <int PREC, int AGGR>
int launch_tc(const float *x, const int *rowptr, const int *col, int N, const float *W, const float *b, float *out,
              int *arg, float *partials, int *n_partials, int *err, cudaStream_t st) {
    const int grid = persistent_grid((N + 7) / 8, 1);
    // >= 116 KB keeps one CTA (and one 512-column TMEM allocation) per SM
    const size_t smem = TcSmem<PREC>::TOTAL > 116 * 1024 ? TcSmem<PREC>::TOTAL : 116 * 1024;
    if (arg) {
        GM_CUDA_TRY(cudaFuncSetAttribute(k_edgeconv_tc<PREC, AGGR, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_edgeconv_tc<PREC, AGGR, true><<<grid, TC_THREADS, smem, st>>>(x, rowptr, col, N, W, b, out, arg, partials, err);
    } else {
        GM_CUDA_TRY(cudaFuncSetAttribute(k_edgeconv_tc<PREC, AGGR, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_edgeconv_tc<PREC, AGGR, false><<<grid, TC_THREADS, smem, st>>>(x, rowptr, col, N, W, b, out, arg, partials, err);
    }
    GM_LAUNCH_CHECK();
    *n_partials = grid;
    return GM_OK;
}

template <int PREC>
int dispatch_aggr(int aggr, const float *x, const int *rowptr, const int *col, int N, const float *W, const float *b,
                  float *out, int *arg, float *partials, int *n_partials, int *err, cudaStream_t st) {
    switch (aggr) {
        case GM_AGGR_MAX: return launch_tc<PREC, GM_AGGR_MAX>(x, rowptr, col, N, W, b, out, arg, partials, n_partials, err, st);
        case GM_AGGR_ADD: return launch_tc<PREC, GM_AGGR_ADD>(x, rowptr, col, N, W, b, out, arg, partials, n_partials, err, st);
        case GM_AGGR_MEAN: return launch_tc<PREC, GM_AGGR_MEAN>(x, rowptr, col, N, W, b, out, arg, partials, n_partials, err, st);
    }
    return GM_ERR_INVALID_ARGUMENT;
}

}  // namespace

// Degrees above TC_LMAX (only possible with max_num_neighbors > 256 or a foreign edge_index) are not tileable.
int edgeconv_tc_forward(const float *x, const int *rowptr, const int *col, int N, const float *W, const float *b,
                        int aggr, int precision, float *out, int *arg, float *partials, int *n_partials,
                        cudaStream_t st) {
    static_assert(sizeof(Book) % 16 == 0, "Book must keep 16-byte alignment across stages");
    if (precision == GM_PREC_TF32)
        return dispatch_aggr<GM_PREC_TF32>(aggr, x, rowptr, col, N, W, b, out, arg, partials, n_partials, nullptr, st);
    if (precision == GM_PREC_BF16)
        return dispatch_aggr<GM_PREC_BF16>(aggr, x, rowptr, col, N, W, b, out, arg, partials, n_partials, nullptr, st);
    return GM_ERR_UNSUPPORTED;
}

}  // namespace gm
