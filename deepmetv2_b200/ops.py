"""Host-side mirror of the three upstream operators the reference's hot path calls, backed by
libgraphmet.so (CUDA, sm_100a) — same names, argument meaning and return types:

* ``radius_graph``  <- torch_cluster.radius_graph   (train.py:48, evaluate.py:88, plt_weight.py:122)
* ``scatter_add``   <- torch_scatter.scatter_add     (model/net.py:8,55-56,79-80,132-133)
* ``knn_graph``     <- torch_cluster.knn_graph      (model/graph_met_network.py:63, dynamic_reduction_network.py:86,95)
* ``build_graph``   the same search without materialising the int64 ``edge_index`` (fused path)

plus ``prepare_inputs`` (train.py:42-46 fused) and ``batch_to_ptr``.
No CPU path: tensors must live on a CUDA device.
"""
from __future__ import annotations

import dataclasses

import torch

from . import _lib
from ._lib import check, lib, ptr, stream


@dataclasses.dataclass
class GraphCSR:
    """CSR by target (rows) with ascending sources (cols) + the transposed CSR for the backward."""
    rowptr: torch.Tensor          # [N+1] int32
    col: torch.Tensor             # [E]   int32
    num_nodes: int
    num_edges: int
    max_degree: int = -1          # before the cap
    num_capped: int = 0
    rowptr_t: torch.Tensor | None = None
    col_t: torch.Tensor | None = None
    capped_flag: torch.Tensor | None = None   # device int32[1] (#capped nodes) when the host never read it back

    @property
    def symmetric(self) -> bool:
        return self.num_capped == 0 and self.capped_flag is None

    def transposed(self):
        """(rowptr_t, col_t): the graph by SOURCE. Equal to (rowptr, col) when no node was capped.  When the cap
        count stayed on the device (``capped_flag``: the sync-free ``edge_capacity`` path) the transposed CSR is
        built by kernels that are no-ops unless the flag is set, and the backward picks the right pair on the device."""
        if self.symmetric:
            return self.rowptr, self.col
        if self.rowptr_t is None:
            N, E = self.num_nodes, self.num_edges
            l = lib()
            ws = torch.empty(l.gm_csr_ws_bytes(N, E), dtype=torch.uint8, device=self.col.device)
            self.rowptr_t = torch.empty(N + 1, dtype=torch.int32, device=self.col.device)
            self.col_t = torch.empty(max(E, 1), dtype=torch.int32, device=self.col.device)
            check(l.gm_csr_transpose(ptr(self.rowptr), ptr(self.col), N, E, ptr(ws), ws.numel(),
                                     ptr(self.rowptr_t), ptr(self.col_t), ptr(self.capped_flag), stream()),
                  "gm_csr_transpose")
        return self.rowptr_t, self.col_t


class EdgeIndex(torch.Tensor):
    """The reference-shaped ``LongTensor[2,E]`` returned by :func:`radius_graph`, carrying the int32
    CSR it was built from so ``GraphMETNetwork.forward`` does not have to re-sort it."""
    __torch_function__ = torch._C._disabled_torch_function_impl

    @staticmethod
    def wrap(t: torch.Tensor, csr: GraphCSR) -> "EdgeIndex":
        e = torch.Tensor._make_subclass(EdgeIndex, t)
        e._gm_csr = csr
        return e


def batch_to_ptr(batch: torch.Tensor, num_events: int) -> torch.Tensor:
    """Sorted int64 ``batch`` [N] -> int32 ``ptr`` [B+1] on the device (no sync)."""
    _lib.require_cuda(batch, "batch")
    batch = batch.contiguous()
    assert batch.dtype == torch.int64
    out = torch.empty(num_events + 1, dtype=torch.int32, device=batch.device)
    check(lib().gm_batch_to_ptr(ptr(batch), batch.numel(), num_events, ptr(out), stream()), "gm_batch_to_ptr")
    return out


def prepare_inputs(x: torch.Tensor):
    """train.py:42-46 in one pass: returns (x_cont view, x_cat_f32 view, etaphi[N,2]).
    phi = atan2f(py, px) is evaluated by the kernel; feed the returned etaphi to an oracle when
    comparing edge sets (libm/ATen atan2 differ from CUDA's in the last ulp)."""
    _lib.require_cuda(x, "x")
    assert x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1 and x.shape[1] >= 11
    etaphi = torch.empty(x.shape[0], 2, dtype=torch.float32, device=x.device)
    check(lib().gm_prepare_etaphi(ptr(x), x.stride(0), x.shape[0], ptr(etaphi), stream()), "gm_prepare_etaphi")
    return x[:, :8], x[:, 8:11], etaphi


def build_graph(pos: torch.Tensor, r: float, ptr_i32: torch.Tensor | None = None, batch: torch.Tensor | None = None,
                loop: bool = False, max_num_neighbors: int = 32, phi_wrap: bool = False,
                max_event_nodes: int | None = None, want_edge_index: bool = False,
                edge_capacity: int | None = None):
    """Cell-binned radius search. Returns GraphCSR (and the int64 edge_index if asked).

    ``edge_capacity``: when given, no host sync happens (E stays on the device, ``num_edges`` is the
    capacity and ``col`` is over-allocated); otherwise E is read back once (as torch_cluster does)."""
    _lib.require_cuda(pos, "x")
    if pos.dim() == 1:
        pos = pos[:, None]
    assert pos.dtype == torch.float32 and pos.stride(1) == 1
    N, dim = pos.shape
    if dim > 2:
        raise NotImplementedError("radius_graph kernels are specialised for 1-D / 2-D (eta, phi) positions")
    dev = pos.device
    if ptr_i32 is None:
        if batch is None:
            ptr_i32 = torch.tensor([0, N], dtype=torch.int32, device=dev)
        else:
            B = int(batch[-1].item()) + 1 if N > 0 else 0   # sorted batch: last element is the max
            ptr_i32 = batch_to_ptr(batch, B)
    B = ptr_i32.numel() - 1
    if max_event_nodes is None:
        max_event_nodes = int((ptr_i32[1:] - ptr_i32[:-1]).max().item()) if B > 0 else 0
    l = lib()
    ws = torch.empty(l.gm_radius_graph_ws_bytes(N, B), dtype=torch.uint8, device=dev)
    rowptr = torch.empty(N + 1, dtype=torch.int32, device=dev)
    stats = torch.zeros(4, dtype=torch.int32, device=dev)
    args = (ptr(pos), pos.stride(0), dim, ptr(ptr_i32), N, B, float(r), int(max_num_neighbors), int(bool(loop)),
            int(bool(phi_wrap)), int(max_event_nodes), ptr(ws), ws.numel())
    check(l.gm_radius_graph_count(*args, ptr(rowptr), ptr(stats), stream()), "gm_radius_graph_count")
    if edge_capacity is None:
        E, max_deg, capped, _flags = stats.tolist()      # one D2H sync, like torch_cluster's masked_select
        cap = E
    else:
        E, max_deg, capped = int(edge_capacity), -1, 0
        cap = E
    col = torch.empty(max(cap, 1), dtype=torch.int32, device=dev)
    ei = torch.empty(2, cap, dtype=torch.int64, device=dev) if want_edge_index else None
    check(l.gm_radius_graph_fill(*args, ptr(rowptr), ptr(col), cap, ptr(ei), ptr(stats), stream()),
          "gm_radius_graph_fill")
    g = GraphCSR(rowptr, col, N, E, max_deg, capped)
    g._stats = stats
    if edge_capacity is not None:       # whether the cap bound is only known on the device: stats[2]
        g.capped_flag = stats[2:3]
    return (g, ei) if want_edge_index else g


@dataclasses.dataclass
class GeoGraph:
    """The radius graph kept implicit (``gm_geo_build``): cell-binned points in internal node order plus the
    per-node cap threshold, inside one opaque device workspace.  ``perm[s]`` = input row of internal row s."""
    ws: torch.Tensor              # uint8 [gm_geo_ws_bytes(N, B)]
    perm: torch.Tensor            # [N] int32
    stats: torch.Tensor           # [4] int32: -1, max degree seen among cap candidates, #capped, flags
    num_nodes: int
    num_events: int


def build_geo(pos: torch.Tensor, r: float, ptr_i32: torch.Tensor, loop: bool = False, max_num_neighbors: int = 32,
              phi_wrap: bool = False, max_event_nodes: int | None = None, out: "GeoGraph | None" = None) -> GeoGraph:
    """Same arguments and edge set as :func:`build_graph`, but no edge list is written (no host sync at all when
    ``max_event_nodes`` is given).  Pass a previous ``out`` of sufficient size to reuse its buffers."""
    _lib.require_cuda(pos, "x")
    if pos.dim() == 1:
        pos = pos[:, None]
    assert pos.dtype == torch.float32 and pos.stride(1) == 1
    N, dim = pos.shape
    if dim > 2:
        raise NotImplementedError("radius_graph kernels are specialised for 1-D / 2-D (eta, phi) positions")
    B = ptr_i32.numel() - 1
    if max_event_nodes is None:
        max_event_nodes = int((ptr_i32[1:] - ptr_i32[:-1]).max().item()) if B > 0 else 0
    l, dev = lib(), pos.device
    need = l.gm_geo_ws_bytes(N, B)
    if out is not None and out.ws.numel() >= need and out.perm.numel() >= N and out.num_events == B \
            and out.num_nodes == N:
        g = out
    else:
        g = GeoGraph(torch.empty(need, dtype=torch.uint8, device=dev),
                     torch.empty(max(N, 1), dtype=torch.int32, device=dev),
                     torch.zeros(4, dtype=torch.int32, device=dev), N, B)
    check(l.gm_geo_build(ptr(pos), pos.stride(0), dim, ptr(ptr_i32), N, B, float(r), int(max_num_neighbors),
                         int(bool(loop)), int(bool(phi_wrap)), int(max_event_nodes), ptr(g.ws), g.ws.numel(),
                         ptr(g.perm), ptr(g.stats), stream()), "gm_geo_build")
    return g


def edgeconv_geo(x_internal: torch.Tensor, geo: GeoGraph, weight: torch.Tensor, bias: torch.Tensor, aggr: str = "max",
                 want_arg: bool = False, precision: str = "fp32"):
    """EdgeConv(Linear(64, 32), aggr) over the implicit graph; ``x_internal`` [N,32] is in internal node order
    (``x[geo.perm]``) and so are the outputs.  ``precision`` "tf32" / "bf16" runs the per-edge contraction on
    tcgen05 (``want_arg`` with ``aggr="max"`` only)."""
    _lib.require_cuda(x_internal, "x")
    x = x_internal.contiguous().float()
    N = x.shape[0]
    l = lib()
    ws = torch.empty(l.gm_edgeconv_ws_bytes(N), dtype=torch.uint8, device=x.device)
    out = torch.empty_like(x)
    arg = torch.empty(N, 32, dtype=torch.int32, device=x.device) if want_arg else None
    check(l.gm_edgeconv_forward_geo(ptr(x), ptr(geo.ws), N, geo.num_events, ptr(weight), ptr(bias), _lib.AGGR[aggr],
                                    _lib.PRECISION[precision], ptr(ws), ws.numel(), ptr(out), ptr(arg), stream()), "gm_edgeconv_forward_geo")
    return (out, arg) if want_arg else out


def radius_graph(x, r, batch=None, loop=False, max_num_neighbors=32, flow="source_to_target", phi_wrap=False):
    """Drop-in for ``torch_cluster.radius_graph``: LongTensor[2,E], row 0 = source j, row 1 = target i,
    sorted by (i, j); at most ``max_num_neighbors`` sources per target (lowest indices kept)."""
    assert flow in ("source_to_target", "target_to_source")
    g, ei = build_graph(x, r, batch=batch, loop=loop, max_num_neighbors=max_num_neighbors, phi_wrap=phi_wrap,
                        want_edge_index=True)
    if flow == "target_to_source":
        return ei.flip(0)
    return EdgeIndex.wrap(ei, g)


def build_knn(x: torch.Tensor, k: int, ptr_i32: torch.Tensor, loop: bool = False) -> GraphCSR:
    """The k-NN search without materialising the int64 ``edge_index`` and WITHOUT a host sync: ``col`` is allocated at
    its upper bound ``N * (k + !loop)`` and ``rowptr[N]`` (= E) stays on the device (``num_edges`` is the capacity).
    What a per-layer dynamic graph (model/graph_met_network.py:63) wants: ``EdgeConv(...)(emb, build_knn(emb, 20, ptr, True))``."""
    _lib.require_cuda(x, "x")
    assert x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1
    N, dim = x.shape
    K = int(k) + (0 if loop else 1)
    if dim > 64 or K > 32:
        raise NotImplementedError("knn_graph kernels cover dim <= 64 and k + (not loop) <= 32")
    dev = x.device
    l = lib()
    ws = torch.empty(l.gm_knn_graph_ws_bytes(N, int(k), int(bool(loop))), dtype=torch.uint8, device=dev)
    rowptr = torch.empty(N + 1, dtype=torch.int32, device=dev)
    cap = max(N * K, 1)
    col = torch.empty(cap, dtype=torch.int32, device=dev)
    check(l.gm_knn_graph_count(ptr(x), x.stride(0), dim, ptr(ptr_i32), N, ptr_i32.numel() - 1, int(k), int(bool(loop)),
                               ptr(ws), ws.numel(), ptr(rowptr), stream()), "gm_knn_graph_count")
    check(l.gm_knn_graph_fill(N, int(k), int(bool(loop)), ptr(ws), ws.numel(), ptr(rowptr), ptr(col), cap, None,
                              stream()), "gm_knn_graph_fill")
    return GraphCSR(rowptr, col, N, cap, -1, num_capped=1)


def knn_graph(x, k, batch=None, loop=False, flow="source_to_target", ptr_i32=None):
    """Drop-in for ``torch_cluster.knn_graph`` (model/graph_met_network.py:63 sketch, dynamic_reduction_network.py:86,95):
    LongTensor[2,E]; per query i its ``k`` nearest nodes j of the same event (``k+1`` searched and the self pair dropped
    when ``loop=False``), ordered by (distance, index); row 0 = j, row 1 = i for ``source_to_target``."""
    assert flow in ("source_to_target", "target_to_source")
    _lib.require_cuda(x, "x")
    if x.dim() == 1:
        x = x[:, None]
    assert x.dtype == torch.float32 and x.stride(1) == 1
    N, dim = x.shape
    dev = x.device
    if ptr_i32 is None:
        if batch is None:
            ptr_i32 = torch.tensor([0, N], dtype=torch.int32, device=dev)
        else:
            B = int(batch[-1].item()) + 1 if N > 0 else 0
            ptr_i32 = batch_to_ptr(batch, B)
    B = ptr_i32.numel() - 1
    if dim > 64 or k + (0 if loop else 1) > 32:
        raise NotImplementedError("knn_graph kernels cover dim <= 64 and k + (not loop) <= 32")
    l = lib()
    ws = torch.empty(l.gm_knn_graph_ws_bytes(N, int(k), int(bool(loop))), dtype=torch.uint8, device=dev)
    rowptr = torch.empty(N + 1, dtype=torch.int32, device=dev)
    check(l.gm_knn_graph_count(ptr(x), x.stride(0), dim, ptr(ptr_i32), N, B, int(k), int(bool(loop)), ptr(ws),
                               ws.numel(), ptr(rowptr), stream()), "gm_knn_graph_count")
    E = int(rowptr[-1].item())                        # one D2H sync, like upstream's `col != -1` mask
    col = torch.empty(max(E, 1), dtype=torch.int32, device=dev)
    ei = torch.empty(2, E, dtype=torch.int64, device=dev)
    check(l.gm_knn_graph_fill(N, int(k), int(bool(loop)), ptr(ws), ws.numel(), ptr(rowptr), ptr(col), E, ptr(ei),
                              stream()), "gm_knn_graph_fill")
    if flow == "target_to_source":
        return ei.flip(0)
    return EdgeIndex.wrap(ei, GraphCSR(rowptr, col, N, E, -1, num_capped=1))     # a kNN graph is not symmetric


def csr_from_edge_index(edge_index: torch.Tensor, num_nodes: int) -> GraphCSR:
    """Any int64 [2,E] edge_index -> GraphCSR (device counting sort; symmetric-ness unknown, so the
    transposed CSR is built on demand for the backward)."""
    csr = getattr(edge_index, "_gm_csr", None)
    if csr is not None and csr.num_nodes == num_nodes:
        return csr
    _lib.require_cuda(edge_index, "edge_index")
    assert edge_index.dtype == torch.int64 and edge_index.dim() == 2 and edge_index.shape[0] == 2
    ei = edge_index.contiguous()
    E = ei.shape[1]
    l = lib()
    ws = torch.empty(l.gm_csr_ws_bytes(num_nodes, E), dtype=torch.uint8, device=ei.device)
    rowptr = torch.empty(num_nodes + 1, dtype=torch.int32, device=ei.device)
    col = torch.empty(max(E, 1), dtype=torch.int32, device=ei.device)
    check(l.gm_edge_index_to_csr(ptr(ei), E, num_nodes, ptr(ws), ws.numel(), ptr(rowptr), ptr(col), stream()),
          "gm_edge_index_to_csr")
    return GraphCSR(rowptr, col, num_nodes, E, -1, num_capped=1)   # treat as possibly asymmetric


def segment_sum(src: torch.Tensor, ptr_i32: torch.Tensor) -> torch.Tensor:
    out = torch.empty(ptr_i32.numel() - 1, dtype=torch.float32, device=src.device)
    check(lib().gm_segment_sum(ptr(src), ptr(ptr_i32), out.numel(), ptr(out), stream()), "gm_segment_sum")
    return out


def scatter_add(src, index, dim=-1, out=None, dim_size=None):
    """Drop-in for ``torch_scatter.scatter_add`` on 1-D float32 CUDA tensors (what model/net.py uses).
    A sorted index (PyG ``batch``) takes the deterministic segmented path; anything else the atomic one.
    Not differentiable — the differentiable MET reduction is ``loss_fn`` / ``met_from_weights``."""
    _lib.require_cuda(src, "src")
    if src.dim() != 1 or index.dim() != 1 or dim not in (-1, 0):
        raise NotImplementedError("scatter_add kernels cover the 1-D case used by the GraphMET path")
    src = src.detach().contiguous().float()
    index = index.contiguous()
    l = lib()
    stats = torch.empty(2, dtype=torch.int64, device=src.device)
    check(l.gm_index_stats(ptr(index), index.numel(), ptr(stats), stream()), "gm_index_stats")
    mx, is_sorted = stats.tolist()
    if out is not None:
        dim_size = out.numel()
    elif dim_size is None:
        dim_size = mx + 1
    if is_sorted and out is None:
        return segment_sum(src, batch_to_ptr(index, dim_size))
    if out is None:
        out = torch.zeros(dim_size, dtype=torch.float32, device=src.device)
    check(l.gm_scatter_add(ptr(src), ptr(index), src.numel(), ptr(out), dim_size, stream()), "gm_scatter_add")
    return out


# ----------------------------------------------------------------------------------------------------------------------
# DynamicReductionNetwork graph operators (row f1): drop-ins for the upstream calls at
# model/dynamic_reduction_network.py:27-30,86-100
# ----------------------------------------------------------------------------------------------------------------------
def to_undirected(edge_index: torch.Tensor, num_nodes: int | None = None):
    """Drop-in for ``torch_geometric.utils.to_undirected``: both directions of every edge, sorted by (row, col),
    duplicates removed.  LongTensor[2,E'] carrying its int32 CSR (``._gm_csr``, plus ``._gm_row``)."""
    _lib.require_cuda(edge_index, "edge_index")
    if num_nodes is None:
        num_nodes = int(edge_index.max().item()) + 1 if edge_index.numel() else 0
    csr = csr_from_edge_index(edge_index, num_nodes)
    N, dev = num_nodes, edge_index.device
    E = edge_index.shape[1]
    l = lib()
    ws = torch.empty(l.gm_undirected_ws_bytes(N, max(E, 1)), dtype=torch.uint8, device=dev)
    cap = max(2 * E, 1)
    rowptr = torch.empty(N + 1, dtype=torch.int32, device=dev)
    col = torch.empty(cap, dtype=torch.int32, device=dev)
    row = torch.empty(cap, dtype=torch.int32, device=dev)
    ei = torch.empty(2, cap, dtype=torch.int64, device=dev)
    check(l.gm_csr_to_undirected(ptr(csr.rowptr), ptr(csr.col), None, None, 0, N, None, ptr(ws), ws.numel(), ptr(rowptr),
                                 ptr(col), ptr(row), cap, ptr(ei), stream()), "gm_csr_to_undirected")
    Eu = int(rowptr[-1].item()) if N else 0       # one D2H sync, like upstream's unique()
    out = EdgeIndex.wrap(ei[:, :Eu].contiguous(), GraphCSR(rowptr, col[:max(Eu, 1)], N, Eu, -1, 0))
    out._gm_row = row[:max(Eu, 1)]
    return out


def normalized_cut_2d(edge_index: torch.Tensor, pos: torch.Tensor) -> torch.Tensor:
    """model/dynamic_reduction_network.py:27-30 (``torch.norm(pos[row] - pos[col])`` then upstream ``normalized_cut``).
    ``edge_index`` must come from :func:`to_undirected` (sorted by row; in-degree == row length)."""
    _lib.require_cuda(pos, "pos")
    csr, row = getattr(edge_index, "_gm_csr", None), getattr(edge_index, "_gm_row", None)
    if csr is None or row is None:
        raise NotImplementedError("normalized_cut_2d takes the edge_index returned by ops.to_undirected")
    pos = pos.contiguous().float()
    E = csr.num_edges
    w = torch.empty(max(E, 1), dtype=torch.float32, device=pos.device)
    check(lib().gm_normalized_cut(ptr(pos), pos.stride(0), pos.shape[1], ptr(csr.rowptr), ptr(csr.col), ptr(row),
                                  csr.num_nodes, E, ptr(w), stream()), "gm_normalized_cut")
    return w[:E]


def graclus(edge_index: torch.Tensor, weight: torch.Tensor | None = None, num_nodes: int | None = None,
            ptr_i32: torch.Tensor | None = None, node_perm: torch.Tensor | None = None, return_partner: bool = False):
    """Drop-in for ``torch_cluster.graclus`` (the CPU algorithm, deterministic visiting order): LongTensor[num_nodes] of
    cluster ids.  ``ptr_i32`` (event boundaries) lets events be matched in parallel (they are independent components);
    ``node_perm`` = visiting order grouped by event (upstream draws a random permutation; default ascending index)."""
    csr = getattr(edge_index, "_gm_csr", None)
    if csr is None:
        raise NotImplementedError("graclus takes the edge_index returned by ops.to_undirected")
    N = csr.num_nodes if num_nodes is None else num_nodes
    dev = edge_index.device
    if ptr_i32 is None:
        ptr_i32 = torch.tensor([0, N], dtype=torch.int32, device=dev)
    men = int((ptr_i32[1:] - ptr_i32[:-1]).max().item())
    cluster = torch.empty(N, dtype=torch.int32, device=dev)
    partner = torch.empty(N, dtype=torch.int32, device=dev)
    visit = None if node_perm is None else node_perm.to(torch.int32).contiguous()
    wt = None if weight is None else weight.contiguous().float()
    check(lib().gm_graclus(ptr(csr.rowptr), ptr(csr.col), ptr(wt), ptr(ptr_i32), N, ptr_i32.numel() - 1, men, ptr(visit),
                           ptr(cluster), ptr(partner), stream()), "gm_graclus")
    return (cluster.long(), partner) if return_partner else cluster.long()


def cluster_max_pool(x: torch.Tensor, cluster: torch.Tensor, partner: torch.Tensor, ptr_i32: torch.Tensor):
    """``torch_geometric.nn.max_pool_x`` for the clusters of :func:`graclus` (size <= 2): returns (x_pooled [M,D],
    ptr_pooled [B+1] int32, consecutive cluster id per node [N] int32).  One D2H sync for M."""
    _lib.require_cuda(x, "x")
    x = x.contiguous().float()
    N, D = x.shape
    B = ptr_i32.numel() - 1
    l, dev = lib(), x.device
    ws = torch.empty(l.gm_undirected_ws_bytes(N, 0), dtype=torch.uint8, device=dev)
    xo = torch.empty(N, D, dtype=torch.float32, device=dev)
    po = torch.empty(B + 1, dtype=torch.int32, device=dev)
    newid = torch.empty(N, dtype=torch.int32, device=dev)
    check(l.gm_cluster_max_pool(ptr(x), D, ptr(cluster.to(torch.int32).contiguous()), ptr(partner), ptr(ptr_i32), N, B, None,
                                ptr(ws), ws.numel(), ptr(xo), ptr(po), ptr(newid), stream()), "gm_cluster_max_pool")
    M = int(po[-1].item())
    return xo[:M], po, newid
