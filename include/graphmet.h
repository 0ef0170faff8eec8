/*
 * graphmet.h — C ABI of libgraphmet.so: the B200 (sm_100a) implementation of the
 * DeepMETv2 GraphMETNetwork hot path.
 *
 * Every entry point is `extern "C"`, takes plain device pointers / sizes / a
 * cudaStream_t (passed as void*), returns an int status (GM_OK == 0, < 0 on error),
 * never throws, never allocates device memory and never synchronises unless stated.
 * The caller (the Python host in deepmetv2_b200/, or any other FFI) owns all
 * buffers.  All kernels are launched on `stream`.
 *
 * Each function cites the reference call it replaces (paths are into the upstream
 * repository DeepMETv2/DeepMETv2).
 *
 * Layouts (fp32 unless stated):
 *   x        [N,11]  PF candidates: pX,pY,pT,eta,d0,dz,mass,puppiWeight | pdgId,charge,fromPV
 *                    (model/data_loader.py:70-77)
 *   ptr      [B+1]   int32 event boundaries into the node axis (batch is sorted)
 *   rowptr   [N+1]   int32 CSR by TARGET node; col[E] int32 = SOURCE node, ascending per row
 *   emb      [N,32]  node embeddings, row-major, hidden_dim = 32 (model/net.py:41-43)
 */
#ifndef GRAPHMET_H_
#define GRAPHMET_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GM_VERSION 100

/* status codes */
#define GM_OK 0
#define GM_ERR_INVALID_ARGUMENT (-1)
#define GM_ERR_CAPACITY (-2)      /* caller-provided buffer too small */
#define GM_ERR_UNSUPPORTED (-3)   /* shape outside what the kernels are specialised for */
#define GM_ERR_CUDA (-4)          /* a CUDA runtime call / launch failed; see gm_last_cuda_error */

/* device-side error flag bits (written with atomicOr into an int the caller provides) */
#define GM_FLAG_CAT_OUT_OF_RANGE 1   /* embedding index outside its table (reference: IndexError) */
#define GM_FLAG_DEGREE_CAPPED 2      /* some node hit max_num_neighbors: CSR is not symmetric */
#define GM_FLAG_EDGE_OVERFLOW 4      /* E exceeded the capacity given to gm_radius_graph_fill */
#define GM_FLAG_EVENT_TOO_LARGE 8     /* an event has more nodes than max_event_nodes promised */

#define GM_HIDDEN 32
#define GM_NUM_PARAM_TENSORS 23
#define GM_NUM_PARAM_FLOATS 6641     /* for continuous_dim = 8, conv_depth = 2 */
#define GM_MAX_CONV_DEPTH 4

/* aggregation of EdgeConv (torch_geometric default for EdgeConv is max) */
#define GM_AGGR_MAX 0
#define GM_AGGR_ADD 1
#define GM_AGGR_MEAN 2

/* arithmetic of the per-edge contraction */
#define GM_PREC_FP32 0      /* fp32 CUDA cores, factored form P_i + aggr_j Y_j (parity mode, default) */
#define GM_PREC_TF32 1      /* tcgen05 kind::tf32 */
#define GM_PREC_BF16 2      /* tcgen05 kind::f16 (bf16 operands, fp32 accumulate in TMEM) */

#define GM_PREC_FP32_DIRECT 4 /* fp32 CUDA cores, per-edge contraction W2 (x_j - x_i) spelled out (cross-check) */

int gm_version(void);
const char *gm_status_string(int status);
int gm_last_cuda_error(void);          /* cudaError_t of the last failing runtime call */
int gm_num_sms(void);                  /* of the current device (cached) */
long long gm_launch_count(int reset);  /* kernels this library launched since the last reset */

/* Optional CUDA-event timing of kernel families on the launching stream (used by bench.py for the
 * live roofline figure). gm_timing_read synchronises on the recorded events, returns the summed
 * milliseconds and the number of timed regions of `slot`, and clears the slot. */
#define GM_TIMER_RADIUS_GRAPH 0
#define GM_TIMER_ENCODE_FWD 1
#define GM_TIMER_EDGECONV_FWD 2
#define GM_TIMER_HEAD_FWD 3
#define GM_TIMER_SCATTER_S_BWD 4
#define GM_TIMER_CONV_NODE_BWD 5
#define GM_TIMER_ENCODE_BWD 6
#define GM_TIMER_BN 7
#define GM_TIMER_NODE_PREP 8
#define GM_TIMER_HEAD_BWD 9
#define GM_TIMER_DRN_NODE 10      /* DRN: inputnet + per-node half of the edge MLP */
#define GM_TIMER_DRN_KNN 11       /* DRN: k-NN search in the 64-d latent space */
#define GM_TIMER_DRN_GRAPH 12     /* DRN: to_undirected */
#define GM_TIMER_DRN_EDGE 13      /* DRN: the per-edge 96->64 contraction kernel alone */
#define GM_TIMER_DRN_POOL 14      /* DRN: BatchNorm merge + apply, normalized cut, graclus, max pool, head */
#define GM_TIMER_DRN_EDGE_BWD 15  /* DRN backward: the per-edge kernel (recompute + BatchNorm / max backward + dW2 + W2^T dz2) */
#define GM_TIMER_DRN_BWD 16       /* DRN backward: everything else (head, sums, mirror, gather, node side, pool, inputnet, reductions) */
#define GM_NUM_TIMERS 17
int gm_timing_enable(int on);
int gm_timing_read(int slot, float *total_ms_out, int *regions_out);

/* Live peak measurements for roofline denominators that MEASURED_PEAKS.json does not carry (both SYNCHRONISE and
 * allocate 4 bytes; measurement helpers, not part of the path): the chip-wide shared-memory read bandwidth (a
 * conflict-free LDS.128 stream on every SM: what physically feeds the cell-staged EdgeConv) and the FFMA stream rate. */
int gm_bench_smem_read(int iters, float *ms_out, double *gbytes_out);
int gm_bench_ffma(int iters, float *ms_out, double *gflop_out);
int gm_bench_ffma2(int iters, float *ms_out, double *gflop_out);   /* packed fma.rn.f32x2 (FFMA2) stream */

/* ---------------------------------------------------------------------------------
 * Parameters: the reference's 23 trainable tensors in construction order
 * (model/graph_met_network.py:15-45; state_dict names in comments) + BN buffers.
 * Gradients use the same struct with the buffers being written.
 * --------------------------------------------------------------------------------- */
typedef struct GmBatchNorm {
    float *weight;               /* [32]  gamma */
    float *bias;                 /* [32]  beta */
    float *running_mean;         /* [32]  (NULL in a gradient struct) */
    float *running_var;          /* [32] */
    int64_t *num_batches_tracked;/* [1]   may be NULL */
} GmBatchNorm;

typedef struct GmParams {
    float *embed_charge;         /* [3,8]   graphnet.embed_charge.weight */
    float *embed_pdgid;          /* [7,8]   graphnet.embed_pdgid.weight */
    float *embed_pv;             /* [8,8]   graphnet.embed_pv.weight */
    float *cont_w, *cont_b;      /* [16,CD],[16]  graphnet.embed_continuous.0.{weight,bias} */
    float *cat_w, *cat_b;        /* [16,24],[16]  graphnet.embed_categorical.0.{weight,bias} */
    float *enc_w, *enc_b;        /* [32,32],[32]  graphnet.encode_all.0.{weight,bias} */
    GmBatchNorm bn_all;          /*               graphnet.bn_all.* */
    float *conv_w[GM_MAX_CONV_DEPTH]; /* [32,64]  graphnet.conv_continuous.l.0.nn.0.weight */
    float *conv_b[GM_MAX_CONV_DEPTH]; /* [32]     graphnet.conv_continuous.l.0.nn.0.bias */
    GmBatchNorm conv_bn[GM_MAX_CONV_DEPTH]; /*    graphnet.conv_continuous.l.1.* */
    float *out1_w, *out1_b;      /* [16,32],[16]  graphnet.output.0.{weight,bias} */
    float *out2_w, *out2_b;      /* [1,16],[1]    graphnet.output.2.{weight,bias} */
    int continuous_dim;          /* CD (8; 7 for the reference's "remove puppi" variant) */
    int conv_depth;              /* 2 for model/net.py Net */
} GmParams;

/* ---------------------------------------------------------------------------------
 * a1. Input preparation — replaces train.py:42-46 / evaluate.py:83-86
 *     (phi = atan2(py,px); etaphi = cat[eta, phi]).  etaphi_out [N,2].
 * --------------------------------------------------------------------------------- */
int gm_prepare_etaphi(const float *x, int x_stride, int N, float *etaphi_out, void *stream);

/* batch[N] (sorted int64, PyG collation) -> ptr[B+1] int32. Replaces the scatter_add+cumsum
 * prologue inside torch_cluster.radius (called from train.py:48). */
int gm_batch_to_ptr(const int64_t *batch, int N, int B, int *ptr_out, void *stream);

/* ---------------------------------------------------------------------------------
 * a2. Radius graph — replaces torch_cluster.radius_graph(etaphi, r, batch, loop,
 *     max_num_neighbors) at train.py:48, evaluate.py:88, plt_weight.py:122.
 *     Cell-binned per-event search; result is bit-identical to the brute force:
 *     edge (j -> i) iff fmaf(dphi,dphi, deta*deta) < (float)(r*r), same event; per target
 *     at most max_num_neighbors sources, lowest source indices first; sorted (i, j).
 *     Two-call protocol: _count fills rowptr (rowptr[N] = E stays on the device; read it
 *     back, or pass an upper bound as capacity), _fill writes col and, optionally, the
 *     reference's int64 edge_index[2,E] (row 0 = source, row 1 = target).
 *     pos: [N, pos_stride] floats with (eta, phi) in columns 0,1 (dim==1: column 0 only).
 *     stats_out (optional, int[4]): {E, max degree before cap, #capped nodes, flags}.
 * --------------------------------------------------------------------------------- */
size_t gm_radius_graph_ws_bytes(int N, int B);
int gm_radius_graph_count(const float *pos, int pos_stride, int dim, const int *ptr, int N, int B,
                          double r, int max_num_neighbors, int loop, int phi_wrap,
                          int max_event_nodes, void *ws, size_t ws_bytes,
                          int *rowptr_out, int *stats_out, void *stream);
int gm_radius_graph_fill(const float *pos, int pos_stride, int dim, const int *ptr, int N, int B,
                         double r, int max_num_neighbors, int loop, int phi_wrap,
                         int max_event_nodes, void *ws, size_t ws_bytes,
                         const int *rowptr, int *col_out, long long col_capacity,
                         int64_t *edge_index64_out /* [2, col_capacity] or NULL */,
                         int *stats_out, void *stream);

/* gm_radius_graph_count + gm_radius_graph_fill in ONE call that allocates nothing. NO COUNTERPART IN THE REFERENCE:
 * added for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it). The workspace is a file-scope static __device__ array of
 * 1,638,400 bytes inside the library (every GPU that loads the library holds it until exit), which bounds the call to
 * N <= GM_STATIC_MAX_NODES and B <= GM_STATIC_MAX_EVENTS; larger sizes return GM_ERR_CAPACITY (and say so on
 * stderr). No cudaMalloc / cudaFree / allocating routine is called; the caller passes every input and output.
 * Results are bit-identical to the two-call protocol: rowptr_out [N+1], col_out [col_capacity] (edges beyond the
 * capacity are dropped and GM_FLAG_EDGE_OVERFLOW is set in stats[3]; E is rowptr_out[N] / stats[0]).
 * ALL calls on a device share the one array and it keeps the last call's contents (the call itself clears what must
 * start at zero): calls that overlap on the GPU (different streams, concurrent threads) RACE — serialise them. */
#define GM_STATIC_MAX_NODES 65536
#define GM_STATIC_MAX_EVENTS 64
int gm_radius_graph_static(const float *pos, int pos_stride, int dim, const int *ptr, int N, int B,
                           double r, int max_num_neighbors, int loop, int phi_wrap,
                           int max_event_nodes, int *rowptr_out, int *col_out, long long col_capacity,
                           int *stats_out, void *stream);

/* Any int64 edge_index[2,E] (e.g. one not produced by this library) -> CSR by target.
 * Device counting sort, sources ascending per row (stable). Needs ws of gm_csr_ws_bytes. */
size_t gm_csr_ws_bytes(int N, long long E);
int gm_edge_index_to_csr(const int64_t *edge_index, long long E, int N, void *ws, size_t ws_bytes,
                         int *rowptr_out, int *col_out, void *stream);

/* ---------------------------------------------------------------------------------
 * f4 / f1. k-nearest-neighbour graph per event in a latent space (dim <= 64) — replaces
 *     torch_cluster.knn_graph(x, k, batch, loop, flow) as sketched at
 *     model/graph_met_network.py:63 (k=20, loop=True on the 32-d embedding) and used by
 *     model/dynamic_reduction_network.py:86,95 (loop=False).
 *     dist = fmaf(df, df, dist) over d in order, df = x_j[d]-x_i[d], fp32; candidates
 *     of the query's own event in ascending index; ties to the lower index; K = k + !loop best
 *     per query ordered by (distance, index); self pair removed afterwards when !loop (so a
 *     query keeps k+1 neighbours when self is not among its K best).  k + !loop <= 32.
 *     Two-call protocol like the radius graph: _count searches, leaves the padded neighbour
 *     table in ws and writes rowptr (rowptr[N] = E on the device); _fill packs col (int32,
 *     CSR by target = query) and, optionally, int64 edge_index[2,capacity] (row 0 = source =
 *     neighbour, row 1 = target = query: flow='source_to_target').
 * --------------------------------------------------------------------------------- */
size_t gm_knn_graph_ws_bytes(int N, int k, int loop);
int gm_knn_graph_count(const float *x, int x_stride, int dim, const int *ptr, int N, int B, int k, int loop,
                       void *ws, size_t ws_bytes, int *rowptr_out, void *stream);
int gm_knn_graph_fill(int N, int k, int loop, const void *ws, size_t ws_bytes, const int *rowptr,
                      int *col_out, long long col_capacity, int64_t *edge_index64_out, void *stream);

/* ---------------------------------------------------------------------------------
 * Whole-network forward — replaces Net.forward / GraphMETNetwork.forward
 * (model/net.py:45-47, model/graph_met_network.py:47-69) given the CSR graph.
 *
 *   x_cont / x_cont_stride : continuous features, row stride in floats (train.py:42 passes the
 *                            strided view data.x[:, :8], stride 11)
 *   x_cat_i64              : [N,3] int64 (train.py:44)  — or NULL and
 *   x_cat_f32 / stride     : the float columns 8..10 of data.x (fused path, no int64 copy)
 *   training               : 1 = batch statistics + running-stat update (model.train()),
 *                            0 = running statistics (model.eval())
 *   save_for_backward      : 1 = also keep the arg-max sources (ws must then be sized with save=1)
 *   apply_sigmoid          : 1 = Net.forward (model/net.py:47), 0 = GraphMETNetwork.forward
 *   node_perm              : int32 [N] or NULL.  When given, the network runs in an INTERNAL node order:
 *                            internal row s is input row node_perm[s]; rowptr/col index internal rows
 *                            (the cell-sorted CSR of gm_radius_graph_*), x_* rows are read at node_perm[s]
 *                            and out is written at node_perm[s], i.e. inputs and outputs stay in the
 *                            caller's order.  NULL = identity.
 *   ws                     : gm_net_ws_bytes(N, depth, save_for_backward) bytes; holds every intermediate the
 *                            backward needs and must stay untouched until gm_net_backward ran
 *   out                    : [N]
 *   err_flags              : int[1], OR-ed with GM_FLAG_* (never cleared by the library)
 * --------------------------------------------------------------------------------- */
size_t gm_net_ws_bytes(int N, int conv_depth, int save_for_backward);
int gm_net_forward(const GmParams *p, const float *x_cont, int x_cont_stride,
                   const int64_t *x_cat_i64, const float *x_cat_f32, int x_cat_stride,
                   const int *rowptr, const int *col, int N, const int *node_perm,
                   int training, int save_for_backward, int apply_sigmoid, int aggr, int precision,
                   float momentum, float eps,
                   void *ws, size_t ws_bytes, float *out, int *err_flags, void *stream);

/* a5 alone — replaces torch_geometric.nn.EdgeConv(nn=Linear(64,32), aggr).forward(x, edge_index)
 * (model/graph_met_network.py:38,65): out_i = aggr_j W.[x_i ; x_j - x_i] + b over the CSR row of i.
 * ws: gm_edgeconv_ws_bytes(N) bytes of scratch (Y = W2 x and P = (W1-W2) x + b; GM_PREC_FP32 only).
 * arg_out (optional, [N,32] int32): arg-max source per (target, channel) for aggr = max. */
size_t gm_edgeconv_ws_bytes(int N);
int gm_edgeconv_forward(const float *x, const int *rowptr, const int *col, int N, const float *W,
                        const float *b, int aggr, int precision, void *ws, size_t ws_bytes, float *out,
                        int *arg_out, void *stream);

/* Backward of gm_net_forward (replaces loss.backward() through the model, train.py:51).
 * grad_out [N] = dL/d out.  Every tensor of `grads` (same layout as GmParams; BN running
 * buffers ignored) is OVERWRITTEN with the gradient.  Deterministic: no atomics; every
 * reduction runs in a fixed order.  The forward must have been run with save_for_backward.
 * (rowptr_t, col_t) is the CSR of the TRANSPOSED graph (by source); for a symmetric graph
 * (no node hit the max_num_neighbors cap) pass (rowptr, col) again.  use_transposed_flag (device
 * int, may be NULL): when given, the kernels use (rowptr_t, col_t) only if *flag != 0 and (rowptr, col)
 * otherwise — lets a sync-free caller decide on the device (pass stats+2 of gm_radius_graph_count).
 * Only aggr = max has a backward so far. */
size_t gm_net_bwd_ws_bytes(int N, int conv_depth);
int gm_net_backward(const GmParams *p, const GmParams *grads, const float *x_cont, int x_cont_stride,
                    const int64_t *x_cat_i64, const float *x_cat_f32, int x_cat_stride,
                    const int *rowptr, const int *col, const int *rowptr_t, const int *col_t,
                    const int *use_transposed_flag, int N, const int *node_perm,
                    int training, int apply_sigmoid, int aggr, float eps,
                    const void *fwd_ws, size_t fwd_ws_bytes, void *bwd_ws, size_t bwd_ws_bytes,
                    const float *out, const float *grad_out, void *stream);

/* ---------------------------------------------------------------------------------
 * Fused path: the radius graph kept IMPLICIT (no edge list is ever written).
 *
 * gm_geo_build replaces the same torch_cluster.radius_graph call (train.py:48, evaluate.py:88) for callers
 * that only feed the graph to the network: it bins the points into the (eta, phi) cell grid, fixes the
 * INTERNAL node order (event, cell, original index) and returns it as node_perm_out [N] (internal row s is
 * input row node_perm[s]), and records per node the max_num_neighbors cap as "largest admissible source
 * index".  The edge j -> i of the reference graph exists iff same event && dist2 < r*r && index(j) <=
 * thr[i] && (loop || j != i) -- exactly the edge set of gm_radius_graph_* (tests compare them bit for bit).
 * gm_net_forward_geo / gm_net_backward_geo / gm_edgeconv_forward_geo are gm_net_forward / gm_net_backward /
 * gm_edgeconv_forward with (geo_ws, B, node_perm) in place of the CSR: the EdgeConv layers stage each cell
 * neighbourhood once into shared memory (TMA bulk copies) and evaluate that predicate there; the backward
 * needs no transposed graph.  precision: GM_PREC_FP32 (factored fp32: forward, backward, arg-max) or GM_PREC_TF32 /
 * GM_PREC_BF16: the per-edge contraction on tcgen05, every consumer warp driving its own TMEM pipeline
 * (csrc/geo_tc.cu); with aggr = max it also records the arg-max sources (save_for_backward / arg_out), so a
 * training step runs on the implicit graph in these precisions too (the backward is the fp32 one, routed through
 * the recorded sources).  geo_ws (gm_geo_ws_bytes(N, B)) must stay
 * untouched between the build and the last forward / backward that uses it.
 * stats_out (optional, int[4]): {-1 (E is not counted), max degree seen among cap candidates, #capped, flags}.
 * x of gm_edgeconv_forward_geo, and arg_out, are in INTERNAL node order.
 * --------------------------------------------------------------------------------- */
size_t gm_geo_ws_bytes(int N, int B);
int gm_geo_build(const float *pos, int pos_stride, int dim, const int *ptr, int N, int B, double r,
                 int max_num_neighbors, int loop, int phi_wrap, int max_event_nodes, void *ws, size_t ws_bytes,
                 int *node_perm_out, int *stats_out, void *stream);
int gm_net_forward_geo(const GmParams *p, const float *x_cont, int x_cont_stride,
                       const int64_t *x_cat_i64, const float *x_cat_f32, int x_cat_stride,
                       const void *geo_ws, int N, int B, const int *node_perm,
                       int training, int save_for_backward, int apply_sigmoid, int aggr, int precision,
                       float momentum, float eps,
                       void *ws, size_t ws_bytes, float *out, int *err_flags, void *stream);
int gm_net_backward_geo(const GmParams *p, const GmParams *grads, const float *x_cont, int x_cont_stride,
                        const int64_t *x_cat_i64, const float *x_cat_f32, int x_cat_stride,
                        const void *geo_ws, int N, int B, const int *node_perm,
                        int training, int apply_sigmoid, int aggr,
                        const void *fwd_ws, size_t fwd_ws_bytes, void *bwd_ws, size_t bwd_ws_bytes,
                        const float *out, const float *grad_out, void *stream);
int gm_edgeconv_forward_geo(const float *x, const void *geo_ws, int N, int B, const float *W, const float *b,
                            int aggr, int precision, void *ws, size_t ws_bytes, float *out, int *arg_out,
                            void *stream);

/* ---------------------------------------------------------------------------------
 * f4. The per-layer DYNAMIC graph sketched at model/graph_met_network.py:63:
 *     emb = emb + BN(EdgeConv(emb, knn_graph(emb, k=20, batch=batch, loop=True)))
 * gm_net_forward_knn / gm_net_backward_knn are gm_net_forward / gm_net_backward with every EdgeConv layer rebuilding
 * its graph from its own input embedding (32-d k-NN per event, csrc/knn.cu; bit-identical to torch_cluster's
 * semantics as restated in oracle/knn_bruteforce.c).  The layer graphs (CSR by target and, when save_for_backward,
 * their transposes -- a k-NN graph is asymmetric) stay in knn_ws (gm_net_knn_ws_bytes(N, conv_depth, k)) between the
 * forward and the backward.  fp32, aggr = max for the backward, nodes in the caller's order, k <= 32.
 * --------------------------------------------------------------------------------- */
size_t gm_net_knn_ws_bytes(int N, int conv_depth, int k);
int gm_net_forward_knn(const GmParams *p, const float *x_cont, int x_cont_stride, const int64_t *x_cat_i64,
                       const float *x_cat_f32, int x_cat_stride, const int *ptr, int N, int B, int k, int training,
                       int save_for_backward, int apply_sigmoid, int aggr, float momentum, float eps, void *ws,
                       size_t ws_bytes, void *knn_ws, size_t knn_ws_bytes, float *out, int *err_flags, void *stream);
int gm_net_backward_knn(const GmParams *p, const GmParams *grads, const float *x_cont, int x_cont_stride,
                        const int64_t *x_cat_i64, const float *x_cat_f32, int x_cat_stride, int N, int k, int training,
                        int apply_sigmoid, int aggr, const void *fwd_ws, size_t fwd_ws_bytes, void *bwd_ws,
                        size_t bwd_ws_bytes, const void *knn_ws, size_t knn_ws_bytes, const float *out,
                        const float *grad_out, void *stream);

/* CSR (by target) -> CSR of the transposed graph (by source), targets ascending per row.
 * ws: gm_csr_ws_bytes(N, E). Needed by the backward only when GM_FLAG_DEGREE_CAPPED was raised;
 * with only_if_flag the kernels are no-ops unless *only_if_flag != 0 (E is then a capacity). */
int gm_csr_transpose(const int *rowptr, const int *col, int N, long long E, void *ws, size_t ws_bytes,
                     int *rowptr_t_out, int *col_t_out, const int *only_if_flag /* device int or NULL */,
                     void *stream);

/* ---------------------------------------------------------------------------------
 * a7. loss_fn — replaces model/net.py:49-62 (two torch_scatter.scatter_add + mean).
 *     px = pred[:,0], py = pred[:,1]; MET_e = sum_{n in e} w_n*(px,py);
 *     loss = 0.5 * mean_e((METx+ty0)^2 + (METy+ty1)^2).  Segmented per-event reduction in a
 *     fixed order (deterministic).  met_out [B,2]; loss_out [1];
 *     grad_w_out (optional) [N] = d loss / d w (so loss.backward() needs no second pass).
 * --------------------------------------------------------------------------------- */
int gm_met_loss(const float *w, const float *pred, int pred_stride, const float *truth,
                int truth_stride, const int *ptr, int N, int B, float *met_out, float *loss_out,
                float *grad_w_out, void *stream);

/* gm_met_loss with an explicit scale on d loss / d w: grad_scale = 1/B reproduces loss_fn's mean (what gm_met_loss
 * passes); grad_scale = 1 yields the gradient of the per-event SUM, which event-sharded data parallelism all-reduces
 * and divides by the GLOBAL event count inside the optimiser kernel (ranks may hold different numbers of events).
 * loss_out is the local mean in either case. */
int gm_met_loss_ex(const float *w, const float *pred, int pred_stride, const float *truth,
                   int truth_stride, const int *ptr, int N, int B, float grad_scale, float *met_out,
                   float *loss_out, float *grad_w_out, void *stream);

/* scatter_add(src, index) for a SORTED index given as ptr (model/net.py:55-56, :132-133). */
int gm_segment_sum(const float *src, const int *ptr, int B, float *out, void *stream);
/* torch_scatter.scatter_add for an arbitrary int64 index (atomicAdd; out must be pre-zeroed /
 * pre-filled by the caller). stats_out int64[2] = {max index (-1 if empty), 1 if sorted}. */
int gm_scatter_add(const float *src, const int64_t *index, long long n, float *out, long long dim_size,
                   void *stream);
int gm_index_stats(const int64_t *index, long long n, int64_t *stats_out, void *stream);

/* a8. resolution metric — replaces model/net.py:92-157 on the device: for each of K MET
 *     flavours (MET = -sum w p first, then truth column pairs (2,3),(4,5),(6,7),(8,9)) writes
 *     u_perp, u_par, response: out [K,3,B]; qt_out [B]. K = 3 if truth has <= 6 columns else 5. */
int gm_resolution(const float *met /*[B,2]*/, const float *truth, int truth_stride, int truth_cols,
                  int B, float *out, float *qt_out, void *stream);

/* f4. The per-epoch summary of evaluate.py:110-156 on the device: events are digitized by qT into n_bins bins of
 *     bin_width (np.digitize over np.arange(0, 400, 10): n_bins = 39, bin_width = 10) and, per MET flavour k and bin,
 *     out[k][0..4][bin] = (q84 - q16)/2 of u_perp, the same of u_perp / mean(R), of u_par, of u_par / mean(R), and
 *     mean(R) -- np.quantile's default linear interpolation; out is [K,5,n_bins+1] with the last column 0 (the
 *     layout of the reference's np.histogram(qT_hist, bins=40, range=(0,400), weights=...)[0]); empty bins are NaN.
 *     vals [K,3,ld]: u_perp, u_par, R of all M events of the epoch (gm_resolution's layout with row pitch ld). */
int gm_resolution_summary(const float *vals, long long ld, const float *qt, int M, int K, float bin_width, int n_bins,
                          float *out, void *stream);

/* ---------------------------------------------------------------------------------
 * a9. Optimiser — replaces torch.optim.AdamW.step (train.py:52,75): one fused kernel over a
 *     flat fp32 buffer. grad_scale multiplies the gradient first (1/world_size after a SUM
 *     allreduce). step is the 1-based step count AFTER this update.
 * --------------------------------------------------------------------------------- */
int gm_adamw_step(float *params, const float *grads, float *exp_avg, float *exp_avg_sq, int n,
                  float lr, float beta1, float beta2, float eps, float weight_decay,
                  int step, float grad_scale, void *stream);

/* gm_adamw_step with (a) an optional DEVICE-side denominator: the gradient is multiplied by
 * grad_scale / max(*grad_denominator_dev, 1) (the global event count that came out of an all-reduce), and (b) an
 * optional status mirror: the kernel copies {step, err_flags_dev[0], graph_stats_dev[3]} into status_host (pinned,
 * device-mapped host memory, int[4]) so the caller can raise on device-side error flags one step late without a
 * copy or a synchronisation. */
int gm_adamw_step_ex(float *params, const float *grads, float *exp_avg, float *exp_avg_sq, int n,
                     float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                     float grad_scale, const float *grad_denominator_dev, const int *err_flags_dev,
                     const int *graph_stats_dev, int *status_host, void *stream);

/* gm_net_backward_geo with the optimiser in its last kernel (single GPU; train.py:51-52 loss.backward() +
 * optimizer.step() as ONE call): the gradient tensors of `grads` must be views that tile opt->grads (the flat
 * gradient buffer, n floats) in the order of opt->params; the last CTA of the final reduction applies
 * gm_adamw_step_ex's update (bit-identical to calling it afterwards) and publishes the status mirror.
 * Contract that saves a memset node per step: the first 64 bytes of bwd_ws (the "last CTA" tickets) must be ZERO
 * the first time a workspace is used; the kernels leave them at zero.  N >= 1. */
typedef struct {
    float *params; const float *grads; float *exp_avg, *exp_avg_sq;
    int n;
    float lr, beta1, beta2, eps, weight_decay;
    int step;                       /* 1-based step count AFTER this update */
    float grad_scale;
    const int *err_flags_dev, *graph_stats_dev;   /* optional status mirror, see gm_adamw_step_ex */
    int *status_host;
} GmAdamW;
int gm_net_backward_geo_adamw(const GmParams *p, const GmParams *grads, const float *x_cont, int x_cont_stride,
                              const int64_t *x_cat_i64, const float *x_cat_f32, int x_cat_stride,
                              const void *geo_ws, int N, int B, const int *node_perm,
                              int training, int apply_sigmoid, int aggr,
                              const void *fwd_ws, size_t fwd_ws_bytes, void *bwd_ws, size_t bwd_ws_bytes,
                              const float *out, const float *grad_out, const GmAdamW *opt, void *stream);

/* ---------------------------------------------------------------------------------
 * e. Event-sharded data parallelism: the ONE collective of the path -- the SUM all-reduce of the flat
 *    gradient (26.6 KB for GraphMETNetwork) -- fused with the optimiser in ONE kernel over NVLink peer memory
 *    (the reference is single-device, train.py:72; this replaces what DistributedDataParallel + optimizer.step()
 *    would do around train.py:51-52).
 *
 *    Every rank owns an exchange buffer (allocated by gm_comm_create with cudaMalloc -- the only allocating call of
 *    the library -- and exported as a 64-byte CUDA IPC handle); gm_comm_connect maps the peers' buffers.
 *    gm_allreduce_adamw_step launches ONE kernel: every CTA stores its slice of
 *        [ grads[0..n) | local_events | local_events * local_loss ]
 *    into slot `rank` of EVERY rank's buffer (remote stores over NVLink / NVSwitch), publishes a per-(rank, CTA) flag
 *    with release semantics, waits for the same slice from every peer, sums the world slices IN RANK ORDER (every
 *    rank computes bit-identical sums, so parameters never diverge), divides by the global event count and applies
 *    AdamW.  No NCCL kernel, no extra launch, no host synchronisation; slots are double-buffered by call parity.
 *    Every rank must call it the same number of times with the same n.  comm == NULL: single process (scale =
 *    1 / local_events).  Waits are bounded (a dead peer traps the kernel after ~20 s instead of hanging the GPU).
 *    global_loss_out (optional, device float[1]): sum_r B_r * loss_r / sum_r B_r.
 *    grads must be the gradient of the per-event SUM (gm_met_loss_ex with grad_scale = 1).
 * --------------------------------------------------------------------------------- */
typedef struct GmComm GmComm;
#define GM_COMM_HANDLE_BYTES 64
#define GM_COMM_MAX_WORLD 16
int gm_comm_create(int n_floats, int rank, int world, GmComm **comm_out, void *ipc_handle_out);
int gm_comm_connect(GmComm *comm, const void *all_handles /* world x GM_COMM_HANDLE_BYTES, rank order */);
int gm_comm_destroy(GmComm *comm);
int gm_allreduce_adamw_step(float *params, const float *grads, float *exp_avg, float *exp_avg_sq, int n,
                            float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                            float local_events, const float *local_loss_dev, GmComm *comm,
                            float *global_loss_out, const int *err_flags_dev, const int *graph_stats_dev,
                            int *status_host, void *stream);

/* ---------------------------------------------------------------------------------
 * f1. DynamicReductionNetwork variant (model/dynamic_reduction_network.py:32-103; wrapper sketched at
 *     model/net.py:12-37: input_dim 11, hidden 64, k 8, output_dim 2, aggr 'max').
 *
 * Graph operators (drop-ins for the upstream calls at dynamic_reduction_network.py:86-102), all on int32 CSR by
 * row with ascending columns; n_dev (optional device int) overrides N with min(N, *n_dev) so a pooled graph
 * whose size only exists on the device needs no host synchronisation.
 *
 *   gm_csr_to_undirected  <- torch_geometric.utils.to_undirected (:86,95): rows i of (rowptr, col) list the
 *       sources j of directed edges j -> i; the result holds every pair in both directions, sorted by (row, col),
 *       duplicates removed.  rowptr_out [N+1] (rowptr_out[N] = E on the device), col_out / row_out [capacity]
 *       (COO row of every entry), optional int64 edge_index[2,capacity] (row 0 = row, row 1 = col); capacity >= 2 x the
 *       number of directed edges, ws >= gm_undirected_ws_bytes(N, directed edges).
 *       Also accepts the padded neighbour table of gm_knn_graph_count instead of a CSR: nbr [N,K], deg [N].
 *   gm_normalized_cut     <- normalized_cut_2d (:27-30) + torch_geometric.utils.normalized_cut:
 *       w_e = ||x[row] - x[col]||_2 * (1/deg[row] + 1/deg[col]).
 *   gm_graclus            <- torch_cluster.graclus (:90,99), the CPU algorithm: per event, nodes visited in `visit`
 *       order (NULL = ascending index; upstream draws a random permutation), an unmatched node takes its
 *       unmatched neighbour of largest weight >= 0 (ties: the later entry of the row); cluster id = the smaller
 *       index of the pair.  cluster_out [N]; partner_out [N] (optional): the other member or -1.
 *   gm_cluster_max_pool   <- torch_geometric.nn.max_pool / max_pool_x (:92,100) for clusters of <= 2 nodes given
 *       as (cluster, partner): consecutive cluster ids in ascending order of the cluster's smallest member,
 *       x_out[c] = max over members, ptr_out [B+1] = pooled event boundaries (ptr_out[B] = M on the device),
 *       newid_out [N] (optional) = consecutive cluster id per input node.  scan_ws: gm_undirected_ws_bytes(N, 0).
 * --------------------------------------------------------------------------------- */
size_t gm_undirected_ws_bytes(int N, long long E_directed_capacity);
int gm_csr_to_undirected(const int *rowptr, const int *col, const int *nbr, const int *deg, int K, int N,
                         const int *n_dev, void *ws, size_t ws_bytes, int *rowptr_out, int *col_out, int *row_out,
                         long long capacity, int64_t *edge_index64_out, void *stream);
int gm_normalized_cut(const float *x, int x_stride, int dim, const int *rowptr, const int *col, const int *row,
                      int N, long long capacity, float *weight_out, void *stream);
int gm_graclus(const int *rowptr, const int *col, const float *weight, const int *ptr, int N, int B,
               int max_event_nodes, const int *visit, int *cluster_out, int *partner_out, void *stream);
int gm_cluster_max_pool(const float *x, int dim, const int *cluster, const int *partner, const int *ptr, int N,
                        int B, const int *n_dev, void *scan_ws, size_t scan_ws_bytes, float *x_out, int *ptr_out,
                        int *newid_out, void *stream);

#define GM_DRN_HIDDEN 64
#define GM_DRN_MIDDLE 96
#define GM_DRN_INPUT 11
typedef struct GmDrnConv {
    float *w1, *b1;              /* [96,128],[96]  drn.edgeconvL.nn.0.{weight,bias}: acts on cat[x_i, x_j - x_i] */
    float *w2, *b2;              /* [64,96],[64]   drn.edgeconvL.nn.2.{weight,bias} */
    float *bn_weight, *bn_bias;  /* [64]           drn.edgeconvL.nn.4.{weight,bias}: BatchNorm1d over the E edge rows */
    float *bn_running_mean, *bn_running_var;       /* [64] */
    int64_t *bn_num_batches_tracked;               /* [1] or NULL */
} GmDrnConv;
typedef struct GmDrnParams {
    float *datanorm;             /* [11]  drn.datanorm */
    float *in_w[3], *in_b[3];    /* drn.inputnet.{0,2,4}: [32,11],[64,32],[64,64] + biases */
    GmDrnConv conv[2];           /* drn.edgeconv1 / drn.edgeconv2 */
    float *out_w[3], *out_b[3];  /* drn.output.{0,2,4}: [64,64],[32,64],[output_dim,32] + biases */
    int output_dim;              /* <= 8 (2 for the wrapper) */
    int k;                       /* neighbours per node (8) */
} GmDrnParams;
/* Device pointers into the workspace of the last gm_drn_forward (for stage-wise parity tests and the backward). */
typedef struct GmDrnView {
    float *x0;                   /* [N,64]   inputnet output */
    int *rowptr1, *col1, *row1;  /* undirected kNN graph of stage 1 (rowptr1[N] = E1) */
    float *x1;                   /* [N,64]   edgeconv1 output */
    float *weight1;              /* [E1]     normalized cut */
    int *cluster1, *partner1;    /* [N] */
    float *x1p;                  /* [M,64]   pooled */
    int *ptr1p;                  /* [B+1]    pooled event boundaries (ptr1p[B] = M) */
    int *rowptr2, *col2, *row2;  /* stage 2 graph over the M pooled nodes (rowptr2[N] = E2) */
    float *x2;                   /* [M,64]   edgeconv2 output */
    float *pooled;               /* [B,64]   global max pool */
    long long edge_capacity;     /* capacity of col / row / weight arrays */
    int *asel1, *asel2;          /* [N,64] (training): source node each EdgeConv's max selected per (node, channel), -1 = no edge */
} GmDrnView;
size_t gm_drn_ws_bytes(int N, int B, int k);
/* Whole forward, no host synchronisation: datanorm, inputnet, then twice (kNN graph in the 64-d latent space ->
 * to_undirected -> EdgeConv(Linear 128->96, ELU, Linear 96->64, ELU, BatchNorm over edges; max) -> [stage 1 only:
 * normalized cut -> graclus -> max_pool]), global max pool, output MLP.  The second normalized_cut / graclus /
 * max_pool_x (:98-100) does not influence the output (global_max_pool of per-cluster maxima is the per-event
 * maximum) and is not run.  training: 1 = BatchNorm batch statistics over the E edge rows + running-stat update.
 * out [B, output_dim] = drn(data); net_out (optional, [B,2], output_dim == 2): softplus(out[:,0]),
 * pi * (2 sigmoid(out[:,1]) - 1) (model/net.py:31-36).  precision: GM_PREC_FP32 or GM_PREC_BF16 / GM_PREC_TF32
 * (the 96 -> 64 per-edge contraction on tcgen05). */
int gm_drn_forward(const GmDrnParams *p, const float *x, int x_stride, const int *ptr, int N, int B,
                   int max_event_nodes, int training, int precision, float momentum, float eps, void *ws,
                   size_t ws_bytes, float *out, float *net_out, GmDrnView *view_out, void *stream);

/* Backward of gm_drn_forward (training = 1, GM_PREC_FP32): grad_out [B, output_dim] = dL/d out.  Every tensor of `grads`
 * (a GmDrnParams whose pointers name the gradient buffers; the BatchNorm running buffers, output_dim and k are ignored;
 * a NULL pointer skips that tensor) is OVERWRITTEN: datanorm, inputnet, both edge MLPs incl. their BatchNorm weight /
 * bias, the output MLP -- 31 tensors.  fwd_ws must be the untouched workspace of the forward on the same inputs.
 * k-NN search, graclus and the pooling indices are constants of the differentiation (as in the reference, where
 * they are integer tensors); max aggregation / max pooling route each gradient to the selected element; BatchNorm
 * over the edge rows is differentiated with batch statistics.  Deterministic (no float atomics). */
/* The wrapper's activations (model/net.py:31-36) fused with a Cartesian MSE on the regressed MET vector:
 * met = softplus(out0), phi = pi (2 sigmoid(out1) - 1), loss = 0.5 mean_e[(met cos phi - t0)^2 + (met sin phi - t1)^2],
 * t = truth[:, 0:2] (genMET px, py).  The reference keeps no loss for this variant (the wrapper is commented out); this
 * one closes the training step.  out [B,2]; net_out (optional) [B,2] = (met, phi); grad_out (optional) [B,2] = dL/d out. */
int gm_drn_met_loss(const float *out, const float *truth, int truth_stride, int B, float *net_out, float *loss_out,
                    float *grad_out, void *stream);
size_t gm_drn_bwd_ws_bytes(int N, int B, int k);
int gm_drn_backward(const GmDrnParams *p, const GmDrnParams *grads, const float *x, int x_stride, const int *ptr, int N,
                    int B, const void *fwd_ws, size_t fwd_ws_bytes, void *bwd_ws, size_t bwd_ws_bytes,
                    const float *grad_out, void *stream);

/* The same exchange as two calls, for callers that put weight-independent work of the NEXT step (its graph build)
 * between them: gm_exchange_post ends step s (pushes this rank's payload to every peer and raises the flags; never
 * waits), gm_exchange_apply_adamw (waits for the peers' payloads of that exchange, rank-ordered sum, global-event-count
 * scaling, AdamW) runs before the next forward.  Every post must be followed by exactly one apply on every rank;
 * parameters end up bit-identical to gm_allreduce_adamw_step. */
int gm_exchange_post(const float *grads, int n, float local_events, const float *local_loss_dev, GmComm *comm, void *stream);
int gm_exchange_apply_adamw(float *params, float *exp_avg, float *exp_avg_sq, int n, float lr, float beta1, float beta2,
                            float eps, float weight_decay, int step, GmComm *comm, float *global_loss_out,
                            const int *err_flags_dev, const int *graph_stats_dev, int *status_host, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* GRAPHMET_H_ */
