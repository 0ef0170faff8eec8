"""ORACLE — TEST INFRASTRUCTURE ONLY.

Stand-ins for the three un-vendored upstream packages (torch_geometric,
torch_cluster, torch_scatter) so that the reference's OWN files
/root/reference/model/graph_met_network.py and /root/reference/model/net.py can
be imported unmodified in this container.  Only the symbols the hot path really
executes are functional (EdgeConv, radius_graph, scatter_add — each backed by
the restatement in graphmet_oracle.py — and, for model/dynamic_reduction_network.py,
knn_graph, to_undirected, normalized_cut, graclus, max_pool, max_pool_x,
global_max_pool from drn_oracle.py); every other imported name is a
placeholder that raises if called.  Used by oracle/make_golden.py only.
"""
import sys
import types

from . import graphmet_oracle as O
from . import drn_oracle as D


def _placeholder(name):
    def _f(*a, **k):
        raise NotImplementedError(f"upstream stub: {name} is not on the GraphMETNetwork hot path")
    _f.__name__ = name
    return _f


class _Cartesian:  # torch_geometric.transforms.Cartesian(cat=False): constructed at import, never used
    def __init__(self, *a, **k):
        pass


def install():
    """Insert the stub packages into sys.modules (idempotent)."""
    if "torch_cluster" in sys.modules and getattr(sys.modules["torch_cluster"], "_is_oracle_stub", False):
        return

    def mod(name, **attrs):
        m = types.ModuleType(name)
        m._is_oracle_stub = True
        for k, v in attrs.items():
            setattr(m, k, v)
        sys.modules[name] = m
        return m

    tc = mod("torch_cluster", radius_graph=O.radius_graph, knn_graph=O.knn_graph)
    ts = mod("torch_scatter", scatter_add=O.scatter_add)
    names_nn = ["GraphConv", "GCNConv", "NNConv", "global_mean_pool", "global_add_pool"]
    drn_nn = dict(graclus=D.graclus, max_pool=D.max_pool, max_pool_x=D.max_pool_x, global_max_pool=D.global_max_pool)
    conv = mod("torch_geometric.nn.conv", EdgeConv=O.EdgeConv,
               **{n: _placeholder(n) for n in ["GraphConv", "GCNConv"]})
    edge_pool = mod("torch_geometric.nn.pool.edge_pool", EdgePooling=_placeholder("EdgePooling"))
    pool = mod("torch_geometric.nn.pool", edge_pool=edge_pool)
    tgnn = mod("torch_geometric.nn", EdgeConv=O.EdgeConv, conv=conv, pool=pool, **drn_nn,
               **{n: _placeholder(n) for n in names_nn})
    undirected = mod("torch_geometric.utils.undirected", to_undirected=D.to_undirected)
    utils = mod("torch_geometric.utils", undirected=undirected,
                normalized_cut=D.normalized_cut,
                remove_self_loops=_placeholder("remove_self_loops"),
                to_undirected=D.to_undirected,
                is_undirected=_placeholder("is_undirected"))
    transforms = mod("torch_geometric.transforms", Cartesian=_Cartesian)
    mod("torch_geometric", nn=tgnn, utils=utils, transforms=transforms)
    return tc, ts
