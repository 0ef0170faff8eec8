"""ORACLE — TEST INFRASTRUCTURE ONLY.

CPU restatement of the reference's DynamicReductionNetwork path
(/root/reference/model/dynamic_reduction_network.py:27-103 and the commented ``Net`` wrapper at
/root/reference/model/net.py:12-37) plus the upstream operators it calls.  Nothing under
``deepmetv2_b200/`` may import this module.

Restated from the reference's own file:
* ``normalized_cut_2d``            dynamic_reduction_network.py:27-30
* ``DynamicReductionNetwork``      dynamic_reduction_network.py:32-103 (module tree / state_dict keys / forward)
* ``DRNNet``                       net.py:12-37 (input_dim 11, hidden 64, k 8, aggr max, softplus / pi(2 sigmoid - 1))

Restated from the published upstream algorithms (torch_geometric 1.7 / torch_cluster 1.5.9, un-vendored,
un-installable here => these pieces are PARITY UNPINNED against upstream binaries; the reference's own file IS
pinned: oracle/make_golden.py imports it unmodified over these stand-ins and tests/test_oracle_drn.py reproduces
the fixtures):
* ``to_undirected``     torch_geometric.utils.to_undirected: cat both directions, coalesce (sort by (row, col), unique)
* ``normalized_cut``    torch_geometric.utils.normalized_cut: w_e * (1/deg[row] + 1/deg[col]), deg = degree(col)
* ``graclus``           torch_cluster.graclus, CPU algorithm (graclus_cpu.cpp): self loops removed, edges grouped by
                        row, nodes visited in ``node_perm`` order (upstream: torch.randperm => RANDOMISED; the identity
                        order used here is one admissible draw), an unmatched node u takes the unmatched neighbour of
                        largest weight (``w >= w_max`` from w_max = 0 in row order: ties to the LATER entry, negative /
                        NaN weights never match), cluster id = min(u, v); unmatched => own id
* ``consecutive_cluster`` / ``max_pool`` / ``max_pool_x`` / ``global_max_pool``   torch_geometric.nn.pool
* ``EdgeConv`` / ``knn_graph``   graphmet_oracle.py (knn_bruteforce.c)
"""
from __future__ import annotations

import math
import types

import numpy as np
import torch
from torch import nn

from . import graphmet_oracle as O


# --------------------------------------------------------------------------- upstream operators
def coalesce(edge_index: torch.Tensor, num_nodes: int) -> torch.Tensor:
    """torch_sparse.coalesce without values: sort by (row, col), drop duplicates."""
    row, col = edge_index[0], edge_index[1]
    key = row * num_nodes + col
    key = torch.unique(key, sorted=True)
    return torch.stack([key // num_nodes, key % num_nodes], dim=0)


def to_undirected(edge_index: torch.Tensor, num_nodes: int | None = None) -> torch.Tensor:
    if num_nodes is None:
        num_nodes = int(edge_index.max()) + 1 if edge_index.numel() else 0
    row, col = edge_index[0], edge_index[1]
    row, col = torch.cat([row, col]), torch.cat([col, row])
    return coalesce(torch.stack([row, col], dim=0), num_nodes)


def degree(index: torch.Tensor, num_nodes: int, dtype=torch.float32) -> torch.Tensor:
    out = torch.zeros(num_nodes, dtype=dtype)
    return out.scatter_add_(0, index, torch.ones(index.numel(), dtype=dtype))


def normalized_cut(edge_index, edge_attr, num_nodes=None):
    row, col = edge_index[0], edge_index[1]
    if num_nodes is None:
        num_nodes = int(edge_index.max()) + 1
    deg = 1.0 / degree(col, num_nodes, edge_attr.dtype)
    deg = deg[row] + deg[col]
    return edge_attr * deg


def normalized_cut_2d(edge_index, pos):
    """dynamic_reduction_network.py:27-30."""
    row, col = edge_index[0], edge_index[1]
    edge_attr = torch.norm(pos[row] - pos[col], p=2, dim=1)
    return normalized_cut(edge_index, edge_attr, num_nodes=pos.size(0))


def graclus(edge_index, weight=None, num_nodes=None, node_perm=None):
    """torch_cluster.graclus (CPU algorithm).  ``node_perm``: visiting order (upstream draws torch.randperm; None =
    identity).  Returns int64 [num_nodes] cluster ids (= min index of the matched pair)."""
    row, col = edge_index[0].numpy(), edge_index[1].numpy()
    if num_nodes is None:
        num_nodes = int(max(row.max(), col.max())) + 1
    mask = row != col
    row, col = row[mask], col[mask]
    w = None if weight is None else weight.detach().numpy()[mask]
    perm = np.argsort(row, kind="stable")
    row, col = row[perm], col[perm]
    if w is not None:
        w = w[perm]
    rowptr = np.zeros(num_nodes + 1, dtype=np.int64)
    np.cumsum(np.bincount(row, minlength=num_nodes), out=rowptr[1:])
    out = np.full(num_nodes, -1, dtype=np.int64)
    order = range(num_nodes) if node_perm is None else [int(v) for v in node_perm]
    for u in order:
        if out[u] >= 0:
            continue
        a, b = rowptr[u], rowptr[u + 1]
        v_max = u
        if w is None:
            for e in range(a, b):
                if out[col[e]] < 0:
                    v_max = int(col[e])
                    break
        else:
            w_max = np.float32(0.0) if w.dtype == np.float32 else 0.0
            for e in range(a, b):
                v = col[e]
                if out[v] >= 0:
                    continue
                if w[e] >= w_max:
                    v_max, w_max = int(v), w[e]
        out[u] = out[v_max] = min(u, v_max)
    return torch.from_numpy(out)


def consecutive_cluster(src: torch.Tensor):
    unique, inv = torch.unique(src, sorted=True, return_inverse=True)
    perm = torch.arange(inv.size(0), dtype=inv.dtype)
    perm = inv.new_empty(unique.size(0)).scatter_(0, inv, perm)
    return inv, perm


def _max_pool_x(cluster, x, size=None):
    size = int(cluster.max()) + 1 if size is None else size
    out = x.new_zeros(size, x.size(1))
    return out.scatter_reduce(0, cluster.view(-1, 1).expand(-1, x.size(1)), x, "amax", include_self=False)


def max_pool_x(cluster, x, batch):
    cluster, perm = consecutive_cluster(cluster)
    return _max_pool_x(cluster, x), batch[perm]


def max_pool(cluster, data):
    """torch_geometric.nn.max_pool: pooled x / batch (the pooled edge_index is rebuilt by the next knn_graph at
    dynamic_reduction_network.py:95, so pool_edge's output is never read; it is still produced for fidelity)."""
    cluster, perm = consecutive_cluster(cluster)
    x = _max_pool_x(cluster, data.x)
    ei = cluster[data.edge_index.view(-1)].view(2, -1)
    ei = ei[:, ei[0] != ei[1]]
    ei = coalesce(ei, x.size(0)) if ei.numel() else ei
    return types.SimpleNamespace(x=x, batch=data.batch[perm], edge_index=ei, edge_attr=None, pos=None)


def global_max_pool(x, batch, size=None):
    size = int(batch.max()) + 1 if size is None else size
    out = x.new_zeros(size, x.size(1))
    return out.scatter_reduce(0, batch.view(-1, 1).expand(-1, x.size(1)), x, "amax", include_self=False)


# --------------------------------------------------------------------------- the reference's own model, restated
class DynamicReductionNetwork(nn.Module):
    """dynamic_reduction_network.py:32-103 with the same attribute names (state_dict keys are the contract).

    ``forward(data, hooks=None)``: ``hooks`` (tests only) may replace the DISCRETE stages so that a float stage can be
    compared on identical graphs: ``hooks["graph"](stage, x, batch) -> edge_index`` and
    ``hooks["cluster"](stage, edge_index, weight, n) -> cluster``; ``hooks["record"]`` (a dict) receives every
    intermediate."""

    def __init__(self, input_dim=5, hidden_dim=64, output_dim=1, k=16, aggr="add",
                 norm=torch.tensor([1. / 500., 1. / 500., 1. / 54., 1 / 25., 1. / 1000.])):
        super().__init__()
        self.datanorm = nn.Parameter(norm)
        self.k = k
        start_width = 2 * hidden_dim
        middle_width = 3 * hidden_dim // 2
        self.inputnet = nn.Sequential(
            nn.Linear(input_dim, hidden_dim // 2), nn.ELU(),
            nn.Linear(hidden_dim // 2, hidden_dim), nn.ELU(),
            nn.Linear(hidden_dim, hidden_dim), nn.ELU())

        def convnn():
            return nn.Sequential(nn.Linear(start_width, middle_width), nn.ELU(),
                                 nn.Linear(middle_width, hidden_dim), nn.ELU(),
                                 nn.BatchNorm1d(num_features=hidden_dim))
        self.edgeconv1 = O.EdgeConv(nn=convnn(), aggr=aggr)
        self.edgeconv2 = O.EdgeConv(nn=convnn(), aggr=aggr)
        self.output = nn.Sequential(nn.Linear(hidden_dim, hidden_dim), nn.ELU(),
                                    nn.Linear(hidden_dim, hidden_dim // 2), nn.ELU(),
                                    nn.Linear(hidden_dim // 2, output_dim))

    def forward(self, data, hooks=None):
        hooks = hooks or {}
        rec = hooks.get("record")

        def note(name, v):
            if rec is not None:
                rec[name] = v.detach().clone() if torch.is_tensor(v) else v

        def graph(stage, x, batch):
            if "graph" in hooks:
                return hooks["graph"](stage, x, batch)
            return to_undirected(O.knn_graph(x, self.k, batch, loop=False, flow="source_to_target"), x.size(0))

        def cluster_of(stage, ei, w, n):
            if "cluster" in hooks:
                return hooks["cluster"](stage, ei, w, n)
            return graclus(ei, w, n)

        x = self.datanorm * data.x                                   # :82
        x = self.inputnet(x)                                         # :83
        note("x0", x)
        ei = graph(1, x, data.batch)                                 # :86
        note("edge_index1", ei)
        x = self.edgeconv1(x, ei)                                    # :87
        note("x1", x)
        weight = normalized_cut_2d(ei, x)                            # :89
        note("weight1", weight)
        cluster = cluster_of(1, ei, weight, x.size(0))               # :90
        note("cluster1", cluster)
        pooled = max_pool(cluster, types.SimpleNamespace(x=x, batch=data.batch, edge_index=ei))   # :92
        x, batch = pooled.x, pooled.batch
        note("x1p", x); note("batch1p", batch)
        ei = graph(2, x, batch)                                      # :95
        note("edge_index2", ei)
        x = self.edgeconv2(x, ei)                                    # :96
        note("x2", x)
        weight = normalized_cut_2d(ei, x)                            # :98
        note("weight2", weight)
        cluster = cluster_of(2, ei, weight, x.size(0))               # :99
        note("cluster2", cluster)
        x, batch = max_pool_x(cluster, x, batch)                     # :100
        x = global_max_pool(x, batch)                                # :102
        note("pooled", x)
        return self.output(x).squeeze(-1)                            # :104


DRN_NORM = [1. / 2950.0, 1. / 2950.0, 1. / 2950.0, 1. / 5.265625, 1. / 143.875, 1. / 589., 1. / 1.2050781, 1. / 211., 1.,
            1. / 7., 1.]


class DRNNet(nn.Module):
    """The commented-out wrapper at model/net.py:12-37."""

    def __init__(self):
        super().__init__()
        self.drn = DynamicReductionNetwork(input_dim=11, hidden_dim=64, k=8, output_dim=2, aggr="max",
                                           norm=torch.tensor(DRN_NORM))

    def forward(self, data, hooks=None):
        output = self.drn(data, hooks)
        met = nn.Softplus()(output[:, 0]).unsqueeze(1)
        metphi = math.pi * (2 * torch.sigmoid(output[:, 1]) - 1).unsqueeze(1)
        return torch.cat((met, metphi), 1)
