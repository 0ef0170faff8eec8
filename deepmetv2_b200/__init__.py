"""deepmetv2_b200 — B200 (sm_100a) implementation of the DeepMETv2 GraphMETNetwork hot path.

Public surface = the reference's own callables (same signatures):
    from deepmetv2_b200.model.net import Net, loss_fn, metrics, resolution, u_perp_par_loss
    from deepmetv2_b200.model.graph_met_network import GraphMETNetwork
    from deepmetv2_b200.ops import radius_graph, scatter_add
CUDA only: every op raises if libgraphmet.so is missing or a tensor is not on a CUDA device.
"""
__version__ = "0.1.0"
