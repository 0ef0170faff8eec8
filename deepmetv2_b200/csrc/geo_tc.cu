// a5 as a REAL per-edge contraction on the 5th-generation tensor cores, fed from the cell-staged IMPLICIT radius graph
// (model/graph_met_network.py:34-39,65; PyG EdgeConv: out_i = aggr_j W [x_i ; x_j - x_i] + b).
//
//   out_i[c] = P_i[c] + aggr_{j in N(i)} sum_k W2[c,k] * round(x_j[k] - x_i[k])      P = W1 x + b (fp32, node side)
//
// Same staging as k_geo_aggr (geo.cu): a producer warp streams the cell neighbourhood of every 32-node work item
// (points + x rows) through a 2-stage shared-memory ring with TMA bulk copies; no edge list exists.  The consumer
// warps differ: each one runs ITS OWN tcgen05 pipeline, with no cross-warp synchronisation at all --
//   * per (target, chunk) it builds the hit list of the edge predicate, then 8 lanes per edge read the staged x_j
//     row (LDS.128), subtract x_i in fp32, round to bf16 / tf32 and append the row to the warp's private K-major
//     operand block (32 edge columns, double-buffered);
//   * the hits of a (target, chunk) are padded to a multiple of 8 columns (max: the last hit again; add / mean: a zero
//     row), so slot runs start on 8-column groups; when 32 columns are full (blocks run across targets and chunks; a
//     partial block is flushed at the end of a segment) one lane issues tcgen05.mma  D[128 x 32] = W2r[128 x 32] * block^T  (W2 replicated 4x along M so that
//     every TMEM lane quarter holds all 32 channels) into the warp's 32 TMEM columns and commits to the warp's
//     mbarrier; the MMA runs while the warp fills the next block;
//   * before the next MMA the warp reads its lane quarter back (tcgen05.ld 32x32b.x32: lane = channel, register =
//     edge column) and folds each 8-column group into the accumulator of the target slot that owns it, in a fixed
//     order -- no atomics, bit-deterministic.
// Tensor work per edge: 2*128*32 flop issued for 2*64*32 useful (W2 replication); the pipe needs ~1 clk per edge per
// SM, TMEM read-back ~2 clk, so the kernel is bound by the shared-memory / issue work of building operand rows.
#include "geo_stage.cuh"
#include "tc_ptx.cuh"

namespace gm {
namespace {

#ifndef GT_CH
#define GT_CH 160                                     // staged neighbourhood rows per pipeline stage
#endif
#ifndef GT_STAGES
#define GT_STAGES 3                                   // 16 consumer warps drain a stage in 2-3 us: one more in flight
#endif
#ifndef GT_NC_BF16
#define GT_NC_BF16 16                                 // consumer warps per CTA (one CTA per SM: 512 TMEM columns)
#endif
#ifndef GT_NC_TF32
#define GT_NC_TF32 8                                  // tf32 operand blocks are twice as large
#endif
constexpr int GT_LIST = GT_CH + 16;                   // build_hit_list pads 16 entries
constexpr int GT_BLK = 32;                            // edge columns per MMA block (MMA N)
constexpr int GT_LBO = 192;                           // bytes between the 16-byte K chunks of an 8-row group: 4 rows x
                                                      // (4 | 8) chunks of a warp store land on distinct banks
using GtStage = StageL<GT_CH, true>;

template <int PREC> struct GtCfg;
template <> struct GtCfg<GM_PREC_TF32> { static constexpr int CHUNKS = 8, KSTEPS = 4, FMT = 2, NC = GT_NC_TF32; };   // 4 tf32 per chunk
template <> struct GtCfg<GM_PREC_BF16> { static constexpr int CHUNKS = 4, KSTEPS = 2, FMT = 1, NC = GT_NC_BF16; };   // 8 bf16 per chunk
constexpr int gt_tmem_cols(int nc) { return nc * GT_BLK <= 32 ? 32 : nc * GT_BLK <= 64 ? 64 : nc * GT_BLK <= 128 ? 128 : nc * GT_BLK <= 256 ? 256 : 512; }

template <int PREC>
struct GtSmem {
    static constexpr int GT_CONSUMERS = GtCfg<PREC>::NC;              // consumer warps (TMEM lane quarter = warp & 3)
    static constexpr int GT_TPW = GEO_TPI / GT_CONSUMERS;             // target slots per warp and item
    static constexpr int THREADS = (GT_CONSUMERS + 1) * 32;           // + 1 producer warp
    static constexpr int TMEM_COLS = gt_tmem_cols(GT_CONSUMERS);
    static_assert(GT_CONSUMERS * GT_BLK <= 512 && GEO_TPI % GT_CONSUMERS == 0 && GT_TPW <= 8, "TMEM columns / slot count");
    static constexpr int SBO = GtCfg<PREC>::CHUNKS * GT_LBO;          // stride of an 8-row group
    static constexpr int A_BYTES = (128 / 8) * SBO;
    static constexpr int B_BYTES = (GT_BLK / 8) * SBO;                // one operand block
    static constexpr size_t OFF_LISTS = (size_t)GT_STAGES * GtStage::BYTES;
    static constexpr size_t OFF_A = OFF_LISTS + (size_t)GT_CONSUMERS * GT_LIST * 4;
    static constexpr size_t OFF_B = OFF_A + A_BYTES;                                   // [warp][2] blocks
    static constexpr size_t OFF_XI = OFF_B + (size_t)GT_CONSUMERS * 2 * B_BYTES;       // [warp][slot][32] f32: x_i
    static constexpr size_t OFF_ACC = OFF_XI + (size_t)GT_CONSUMERS * GT_TPW * 128;    // [warp][slot][32] f32
    static constexpr size_t OFF_SLOT = OFF_ACC + (size_t)GT_CONSUMERS * GT_TPW * 128;  // [warp][slot] (eta, phi, thr, excl)
    static constexpr size_t OFF_DEG = OFF_SLOT + (size_t)GT_CONSUMERS * GT_TPW * 16;   // [warp][slot] i32
    static constexpr size_t OFF_GSLOT = OFF_DEG + (size_t)GT_CONSUMERS * GT_TPW * 4;   // [warp][2][4] u8: slot of a group
    static constexpr size_t OFF_CNODE = (OFF_GSLOT + (size_t)GT_CONSUMERS * 8 + 15) / 16 * 16;  // [warp][2][32] i32: source node of a column (arg-max variant)
    static constexpr size_t OFF_ARG = OFF_CNODE + (size_t)GT_CONSUMERS * 2 * GT_BLK * 4;        // [warp][slot][32] i32: running arg-max
    static constexpr size_t OFF_BAR = (OFF_ARG + (size_t)GT_CONSUMERS * GT_TPW * 128 + 7) / 8 * 8;     // full[2] empty[2] mma[warp]
    static constexpr size_t OFF_TMEM = OFF_BAR + (size_t)(2 * GT_STAGES + GT_CONSUMERS) * 8;
    static constexpr size_t TOTAL = OFF_TMEM + 16;
    static_assert(OFF_A % 128 == 0 && OFF_B % 16 == 0 && OFF_XI % 16 == 0 && OFF_SLOT % 16 == 0, "operand alignment");
};

// K-major, no-swizzle operand layout: 8-row x 16-byte core matrices; chunk stride LBO, 8-row group stride SBO
template <int PREC>
__device__ __forceinline__ uint64_t gt_smem_desc(uint32_t addr) {
    uint64_t d = (uint64_t)((addr >> 4) & 0x3FFF);
    d |= (uint64_t)(GT_LBO >> 4) << 16;
    d |= (uint64_t)(GtSmem<PREC>::SBO >> 4) << 32;
    d |= 1ull << 46;                          // descriptor version (sm_100)
    return d;                                 // layout_type = 0 (SWIZZLE_NONE), base_offset = 0
}
template <int PREC>
__device__ __forceinline__ uint32_t gt_instr_desc(int n) {
    return (1u << 4) | ((uint32_t)GtCfg<PREC>::FMT << 7) | ((uint32_t)GtCfg<PREC>::FMT << 10) |
           ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);   // F32 accum, K-major A and B, M = 128
}
__device__ __forceinline__ void gt_sts64(uint32_t addr, uint32_t a, uint32_t b) {
    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ void gt_sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// ARG (max only): also the arg-max SOURCE node per (target, channel), so that tf32 / bf16 TRAINING runs on the implicit
// graph: the source node of every operand column is recorded when the column is filled (the stage may be gone by the
// time the block comes back from the tensor core) and the epilogue keeps the first maximum in column order.
template <int PREC, int AGGR, bool ARG = false>
__global__ void __launch_bounds__(GtSmem<PREC>::THREADS, 1)
k_geo_edgeconv_tc(const float *__restrict__ X, const float *__restrict__ P, const float *__restrict__ W, GeoView v,
                  int N, float *__restrict__ out, int *__restrict__ arg) {
    using S = GtSmem<PREC>;
    using C = GtCfg<PREC>;
    constexpr bool TF32 = PREC == GM_PREC_TF32;
    constexpr int GT_CONSUMERS = S::GT_CONSUMERS, GT_TPW = S::GT_TPW, GT_THREADS = S::THREADS, GT_TMEM_COLS = S::TMEM_COLS;
    extern __shared__ __align__(128) unsigned char gs[];
    const uint32_t full0 = smem_u32(gs + S::OFF_BAR), empty0 = full0 + 8u * GT_STAGES, mma0 = empty0 + 8u * GT_STAGES;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const GeoMeta m = *v.meta;
    const float IDENT = AGGR == GM_AGGR_MAX ? -INFINITY : 0.f;

    // ---- one-time setup: barriers, TMEM, the A operand (W2 replicated 4x along M, rounded to the MMA input type)
    if (tid == 0) {
        for (int s = 0; s < GT_STAGES; ++s) { mbar_init(full0 + 8u * s, 1); mbar_init(empty0 + 8u * s, GT_CONSUMERS); }
        for (int w = 0; w < GT_CONSUMERS; ++w) mbar_init(mma0 + 8u * w, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) tc5_alloc(smem_u32(gs + S::OFF_TMEM), GT_TMEM_COLS);
    const uint32_t a_base = smem_u32(gs + S::OFF_A);
    for (int t = tid; t < 128 * C::CHUNKS; t += GT_THREADS) {
        const int row = t / C::CHUNKS, c = t - row * C::CHUNKS;
        const float *wrow = W + (row & 31) * CONV_IN + H;
        const uint32_t addr = a_base + (uint32_t)((row >> 3) * S::SBO + c * GT_LBO + (row & 7) * 16);
        if constexpr (TF32) {
            const float4 w0 = *reinterpret_cast<const float4 *>(wrow + 4 * c);
            gt_sts128(addr, tc5_to_tf32(w0.x), tc5_to_tf32(w0.y), tc5_to_tf32(w0.z), tc5_to_tf32(w0.w));
        } else {
            const float4 w0 = *reinterpret_cast<const float4 *>(wrow + 8 * c);
            const float4 w1 = *reinterpret_cast<const float4 *>(wrow + 8 * c + 4);
            gt_sts128(addr, tc5_pack_bf16(w0.x, w0.y), tc5_pack_bf16(w0.z, w0.w), tc5_pack_bf16(w1.x, w1.y),
                      tc5_pack_bf16(w1.z, w1.w));
        }
    }
    tc5_fence_async_smem();
    tc5_fence_before();
    __syncthreads();
    tc5_fence_after();
    pdl_sync();                                          // X, P and the tickets are the previous kernels' outputs
    const uint32_t tmem_base = *reinterpret_cast<const uint32_t *>(gs + S::OFF_TMEM);

    if (warp == GT_CONSUMERS) {
        // ================================================================== producer (tickets: no BatchNorm partials here)
        geo_produce<GtStage, true, GT_STAGES>(gs, full0, empty0, v, m, X, N, true, lane, 200);
    } else {
        // ================================================================== consumers
        const int w = warp;
        unsigned *list = reinterpret_cast<unsigned *>(gs + S::OFF_LISTS) + w * GT_LIST;
        float *s_xi = reinterpret_cast<float *>(gs + S::OFF_XI) + w * GT_TPW * 32;
        float *s_acc = reinterpret_cast<float *>(gs + S::OFF_ACC) + w * GT_TPW * 32;
        float4 *s_slot = reinterpret_cast<float4 *>(gs + S::OFF_SLOT) + w * GT_TPW;
        int *s_deg = reinterpret_cast<int *>(gs + S::OFF_DEG) + w * GT_TPW;
        unsigned char *s_gslot = gs + S::OFF_GSLOT + w * 8;
        int *s_cnode = reinterpret_cast<int *>(gs + S::OFF_CNODE) + w * 2 * GT_BLK;
        int *s_arg = reinterpret_cast<int *>(gs + S::OFF_ARG) + w * GT_TPW * 32;
        const uint32_t b_base = smem_u32(gs + S::OFF_B) + (uint32_t)(w * 2 * S::B_BYTES);
        const uint32_t mbar = mma0 + 8u * w;
        const uint32_t d_tmem = tmem_base + (uint32_t)(w * GT_BLK);                 // all 128 lanes, this warp's columns
        const uint32_t ld_tmem = d_tmem + ((uint32_t)((w & 3) * 32) << 16);         // this warp's lane quarter
        const uint32_t idesc = gt_instr_desc<PREC>(GT_BLK);
        const uint64_t adesc0 = gt_smem_desc<PREC>(a_base), bdesc0 = gt_smem_desc<PREC>(b_base);
        const int r4 = lane >> 3, c4 = lane & 7;                                    // 8 lanes per edge row, 4 channels per lane
        const uint32_t lane_off = TF32 ? (uint32_t)(c4 * GT_LBO) : (uint32_t)((c4 >> 1) * GT_LBO + (c4 & 1) * 8);
        int fill = 0, buf = 0, pending = -1, nv_pending = 0;
        uint32_t mma_parity = 0;

        // read the finished block back (lane = channel, register = edge column) and fold its columns into the slots
        auto epilogue = [&]() {
            mbar_wait(mbar, mma_parity, 32);
            mma_parity ^= 1u;
            tc5_fence_after();
            uint32_t r[32];
            tc5_ld32(ld_tmem, r);
            const unsigned slots = *reinterpret_cast<const unsigned *>(s_gslot + pending * 4);
            float run = IDENT;                                                      // groups of one slot are merged in registers
            int runcol = 0;
#pragma unroll
            for (int g = 0; g < GT_BLK / 8; ++g) {
                if (g * 8 < nv_pending) {                                           // warp-uniform
                    float t[8];
#pragma unroll
                    for (int c = 0; c < 8; ++c) t[c] = __uint_as_float(r[8 * g + c]);
                    if constexpr (ARG) {
#pragma unroll
                        for (int c = 0; c < 8; ++c)
                            if (t[c] > run) { run = t[c]; runcol = 8 * g + c; }     // strict: the first maximum wins
                    } else if (AGGR == GM_AGGR_MAX)
                        run = fmaxf(fmaxf(fmaxf(run, t[0]), fmaxf(t[1], t[2])), fmaxf(fmaxf(t[3], t[4]), fmaxf(t[5], fmaxf(t[6], t[7]))));
                    else
                        run += ((t[0] + t[1]) + (t[2] + t[3])) + ((t[4] + t[5]) + (t[6] + t[7]));
                    const unsigned sg = (slots >> (8 * g)) & 0xffu;
                    const bool close = (g + 1) * 8 >= nv_pending || g == GT_BLK / 8 - 1 || ((slots >> (8 * g + 8)) & 0xffu) != sg;
                    if (close) {                                                    // warp-uniform
                        float *dst = s_acc + (int)sg * 32 + lane;
                        if constexpr (ARG) {
                            if (run > *dst) { *dst = run; s_arg[(int)sg * 32 + lane] = s_cnode[pending * GT_BLK + runcol]; }
                        } else {
                            *dst = AGGR == GM_AGGR_MAX ? fmaxf(*dst, run) : *dst + run;
                        }
                        run = IDENT;
                    }
                }
            }
            tc5_fence_before();
            pending = -1;
        };
        // hand the current block (columns 0..fill-1 valid) to the tensor core
        auto flush = [&]() {
            if (pending >= 0) epilogue();                                           // frees the warp's TMEM columns
            tc5_fence_async_smem();                                                 // operand stores -> async proxy
            __syncwarp();
            if (lane == 0) {
                tc5_fence_after();
                const uint64_t bdesc = bdesc0 + (uint64_t)((buf * S::B_BYTES) >> 4);   // address field only: no carry out
#pragma unroll
                for (int ks = 0; ks < C::KSTEPS; ++ks)
                    tc5_mma<TF32>(d_tmem, adesc0 + (uint64_t)((ks * 2 * GT_LBO) >> 4), bdesc + (uint64_t)((ks * 2 * GT_LBO) >> 4),
                                  idesc, ks > 0 ? 1u : 0u);
                tc5_commit(mbar);
            }
            __syncwarp();
            pending = buf; nv_pending = fill; buf ^= 1; fill = 0;
        };

        for (uint32_t gc = 0;; ++gc) {
            const int st = gc % GT_STAGES;
            const unsigned char *stage = gs + st * GtStage::BYTES;
            mbar_wait(full0 + 8u * st, (gc / GT_STAGES) & 1u);
            const ChunkDesc *d = reinterpret_cast<const ChunkDesc *>(stage + GtStage::OFF_DESC);
            const int flags = d->flags;
            if (flags & AG_END) break;
            const int a = d->a, b = d->b, rows = d->rows, n0 = d->n0;
            const float4 *s_pts = reinterpret_cast<const float4 *>(stage + GtStage::OFF_PTS);
            if (flags & AG_FIRST) {
                const float4 *s_tgt = reinterpret_cast<const float4 *>(stage + GtStage::OFF_TGT);
                const int *s_thr = reinterpret_cast<const int *>(stage + GtStage::OFF_THR);
                const float *s_trows = reinterpret_cast<const float *>(stage + GtStage::OFF_TROWS);
                if (lane < GT_TPW) {
                    const int t = a + w + lane * GT_CONSUMERS;
                    if (t < b) {
                        const float4 pt = s_tgt[t];
                        s_slot[lane] = make_float4(pt.x, pt.y, __int_as_float(s_thr[t]), m.loop ? __int_as_float(-1) : pt.z);
                        s_deg[lane] = 0;
                    }
                }
#pragma unroll
                for (int u = 0; u < GT_TPW; ++u) {
                    const int t = a + w + u * GT_CONSUMERS;
                    s_acc[u * 32 + lane] = IDENT;
                    if constexpr (ARG) s_arg[u * 32 + lane] = -1;
                    if (t < b) s_xi[u * 32 + lane] = s_trows[t * 32 + lane];
                }
                __syncwarp();
            }
#pragma unroll 1
            for (int u = 0; u < GT_TPW; ++u) {
                if (a + w + u * GT_CONSUMERS >= b) break;                           // warp-uniform
                const float4 sl = s_slot[u];
                const int nh = build_hit_list<GT_CH>(s_pts, rows, sl.x, sl.y, __float_as_int(sl.z), __float_as_int(sl.w),
                                                     m, lane, list);
                if (lane == 0) s_deg[u] += nh;
                const float4 xi = *reinterpret_cast<const float4 *>(s_xi + u * 32 + c4 * 4);
                if (nh > 0) {
                    // pad the list to whole 8-column groups: max -> the last hit again; add / mean -> x_i itself (a zero row)
                    const unsigned pad = AGGR == GM_AGGR_MAX ? list[nh - 1]
                                                             : (unsigned)(smem_u32(s_xi + u * 32) - smem_u32(stage));
                    __syncwarp();
                    if (lane < 8) list[nh + lane] = pad;
                    __syncwarp();
                    const uint32_t src = smem_u32(stage) + (uint32_t)(c4 * 16);
                    const uint32_t dstl = b_base + lane_off + (uint32_t)(r4 * 16);
                    // the padded hits go into the block in runs bounded by the block end: the inner loop over 8-column
                    // groups carries no block bookkeeping (two running pointers), the compiler unrolls it by two
                    auto lds4 = [](float4 &q, uint32_t a) {
                        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(q.x), "=f"(q.y), "=f"(q.z), "=f"(q.w) : "r"(a));
                    };
                    const int ngroups = (nh + 7) >> 3;
                    const unsigned *lp = list + r4;
                    int kg = 0;
                    while (kg < ngroups) {                                          // warp-uniform
                        const int ng = min(ngroups - kg, (GT_BLK - fill) >> 3);
                        if (lane < ng) s_gslot[buf * 4 + (fill >> 3) + lane] = (unsigned char)u;
                        uint32_t ad = dstl + (uint32_t)(buf * S::B_BYTES + (fill >> 3) * S::SBO);
                        int *cn = s_cnode + buf * GT_BLK + fill + r4;
#pragma unroll 2
                        for (int g = 0; g < ng; ++g, lp += 8, ad += S::SBO, cn += 8) {
                            float4 x0, x1;
                            lds4(x0, src + lp[0]);
                            lds4(x1, src + lp[4]);
                            if constexpr (ARG) {
                                if (c4 == 0) {                                      // one lane per staged row: column -> node
                                    cn[0] = node_of_row(d, (int)(lp[0] >> 7));
                                    cn[4] = node_of_row(d, (int)(lp[4] >> 7));
                                }
                            }
                            if constexpr (TF32) {
                                gt_sts128(ad, tc5_to_tf32(x0.x - xi.x), tc5_to_tf32(x0.y - xi.y), tc5_to_tf32(x0.z - xi.z), tc5_to_tf32(x0.w - xi.w));
                                gt_sts128(ad + 64, tc5_to_tf32(x1.x - xi.x), tc5_to_tf32(x1.y - xi.y), tc5_to_tf32(x1.z - xi.z), tc5_to_tf32(x1.w - xi.w));
                            } else {
                                gt_sts64(ad, tc5_pack_bf16(x0.x - xi.x, x0.y - xi.y), tc5_pack_bf16(x0.z - xi.z, x0.w - xi.w));
                                gt_sts64(ad + 64, tc5_pack_bf16(x1.x - xi.x, x1.y - xi.y), tc5_pack_bf16(x1.z - xi.z, x1.w - xi.w));
                            }
                        }
                        fill += 8 * ng; kg += ng;
                        if (fill == GT_BLK) flush();
                    }
                }
                __syncwarp();                                   // the hit list is reused by the next target
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty0 + 8u * st);       // this warp is done with the stage
            if (flags & AG_LAST) {
                float p[GT_TPW];
#pragma unroll
                for (int u = 0; u < GT_TPW; ++u) {
                    const int t = a + w + u * GT_CONSUMERS;
                    p[u] = t < b ? __ldg(P + (size_t)(n0 + t) * H + lane) : 0.f;
                }
                if (fill > 0) flush();
                if (pending >= 0) epilogue();
                __syncwarp();
#pragma unroll
                for (int u = 0; u < GT_TPW; ++u) {
                    const int t = a + w + u * GT_CONSUMERS;
                    if (t >= b) continue;
                    const int deg = s_deg[u];
                    const float acc = s_acc[u * 32 + lane];
                    float o;
                    if (deg == 0) o = 0.f;                                          // PyG: empty rows aggregate to 0
                    else if (AGGR == GM_AGGR_MAX) o = p[u] + acc;
                    else if (AGGR == GM_AGGR_ADD) o = fmaf((float)deg, p[u], acc);
                    else o = fmaf(acc, 1.f / (float)deg, p[u]);
                    out[(size_t)(n0 + t) * H + lane] = o;
                    if constexpr (ARG) arg[(size_t)(n0 + t) * H + lane] = deg == 0 ? -1 : s_arg[u * 32 + lane];
                }
                __syncwarp();
            }
        }
    }

    // ---- teardown: TMEM back, last CTA out re-arms the ticket counter
    tc5_fence_before();
    __syncthreads();
    tc5_fence_after();
    if (warp == 0) tc5_dealloc(tmem_base, GT_TMEM_COLS);
    if (tid == 0) {
        __threadfence();
        if (atomicAdd(v.sched + 1, 1) == (int)gridDim.x - 1) { v.sched[0] = 0; v.sched[1] = 0; __threadfence(); }
    }
}

template <int PREC, int AGGR, bool ARG = false>
int launch_geo_tc(const float *X, const float *P, const float *W, const GeoView &v, int N, float *out, int *arg,
                  cudaStream_t st) {
    static int occ = 0;
    if (!occ) {
        GM_CUDA_TRY(cudaFuncSetAttribute(k_geo_edgeconv_tc<PREC, AGGR, ARG>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)GtSmem<PREC>::TOTAL));
        GM_CUDA_TRY(cudaFuncSetAttribute(k_geo_edgeconv_tc<PREC, AGGR, ARG>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                         (int)cudaSharedmemCarveoutMaxShared));
        occ = min(resident_ctas(k_geo_edgeconv_tc<PREC, AGGR, ARG>, GtSmem<PREC>::THREADS, GtSmem<PREC>::TOTAL, 1), 512 / GtSmem<PREC>::TMEM_COLS);
        // work is taken from a ticket counter, so a CTA that is not resident at once only finds the queue empty:
        // size the grid by what shared memory allows even if the occupancy query is pessimistic
        occ = max(occ, min((int)(227 * 1024 / (GtSmem<PREC>::TOTAL + 1024)), 512 / GtSmem<PREC>::TMEM_COLS));
    }
    const int grid = persistent_grid((N + GEO_TPI - 1) / GEO_TPI, occ);
    GM_LAUNCH_PDL((k_geo_edgeconv_tc<PREC, AGGR, ARG>), grid, GtSmem<PREC>::THREADS, GtSmem<PREC>::TOTAL, st, X, P, W, v, N, out, arg);
    GM_LAUNCH_CHECK();
    return GM_OK;
}

template <int PREC>
int dispatch_geo_tc(int aggr, const float *X, const float *P, const float *W, const GeoView &v, int N, float *out,
                    int *arg, cudaStream_t st) {
    if (arg) {
        if (aggr != GM_AGGR_MAX) return GM_ERR_UNSUPPORTED;
        return launch_geo_tc<PREC, GM_AGGR_MAX, true>(X, P, W, v, N, out, arg, st);
    }
    switch (aggr) {
        case GM_AGGR_MAX: return launch_geo_tc<PREC, GM_AGGR_MAX>(X, P, W, v, N, out, nullptr, st);
        case GM_AGGR_ADD: return launch_geo_tc<PREC, GM_AGGR_ADD>(X, P, W, v, N, out, nullptr, st);
        case GM_AGGR_MEAN: return launch_geo_tc<PREC, GM_AGGR_MEAN>(X, P, W, v, N, out, nullptr, st);
    }
    return GM_ERR_INVALID_ARGUMENT;
}

}  // namespace

// X: layer input [N,32] in internal order; P = W1 x + b [N,32]; W: the EdgeConv weight [32,64] (W2 = columns 32..63)
// arg (optional, max only): [N,32] arg-max source node per (target, channel), internal order
int geo_edgeconv_tc(int aggr, int precision, const float *X, const float *P, const float *W, const void *geo_ws, int N,
                    int B, float *out, int *arg, cudaStream_t st) {
    if (!geo_ws || B < 1) return GM_ERR_INVALID_ARGUMENT;
    const GeoView v = view_of(geo_ws, N, B);
    if (precision == GM_PREC_TF32) return dispatch_geo_tc<GM_PREC_TF32>(aggr, X, P, W, v, N, out, arg, st);
    if (precision == GM_PREC_BF16) return dispatch_geo_tc<GM_PREC_BF16>(aggr, X, P, W, v, N, out, arg, st);
    return GM_ERR_UNSUPPORTED;
}

}  // namespace gm
