"""Pins oracle/knn_bruteforce.c (the restatement of torch_cluster.knn_graph; model/graph_met_network.py:63 sketch,
model/dynamic_reduction_network.py:86,95) with cases whose answer is known independently of fp rounding: integer
coordinates (every distance exact in fp32, so a stable sort over (distance, index) is the ground truth)."""
import torch

from oracle import graphmet_oracle as O


def _stable_knn(x, k, batch, loop):
    """Independent statement: fp64 distances of integer points are exact; stable sort => ties to the lower index."""
    src, dst = [], []
    K = k if loop else k + 1
    for e in torch.unique(batch).tolist():
        idx = (batch == e).nonzero().flatten()
        p = x[idx].double()
        d = ((p[:, None, :] - p[None, :, :]) ** 2).sum(-1)
        order = torch.sort(d, dim=1, stable=True).indices[:, :K]
        for qi in range(idx.numel()):
            nb = idx[order[qi]]
            if not loop:
                nb = nb[nb != idx[qi]]
            src.append(nb); dst.append(idx[qi].repeat(nb.numel()))
    return torch.stack([torch.cat(src), torch.cat(dst)], 0)


def test_knn_oracle_integer_points_match_stable_sort():
    g = torch.Generator().manual_seed(3)
    for dim, k, loop in [(2, 4, False), (3, 8, True), (32, 20, True), (64, 8, False), (5, 31, False)]:
        sizes = [40, 1, 7, 100, 2]
        batch = torch.repeat_interleave(torch.arange(len(sizes)), torch.tensor(sizes))
        x = torch.randint(-3, 4, (sum(sizes), dim), generator=g).float()      # many ties and exact duplicates
        got = O.knn_graph(x, k, batch, loop=loop)
        want = _stable_knn(x, k, batch, loop)
        assert torch.equal(got, want), (dim, k, loop)


def test_knn_oracle_duplicates_keep_k_plus_one_when_self_is_not_among_the_best():
    """upstream masks row != col AFTER asking for k+1: with > k duplicates at lower indices the self pair never enters
    the list, and the node keeps k+1 neighbours."""
    x = torch.zeros(6, 2)
    ei = O.knn_graph(x, 2, None, loop=False)
    deg = torch.bincount(ei[1], minlength=6)
    assert deg.tolist() == [2, 2, 2, 3, 3, 3]
    assert ei[0, ei[1] == 5].tolist() == [0, 1, 2]


def test_knn_oracle_small_events_flow_and_order():
    x = torch.tensor([[0.], [1.], [3.], [10.], [11.]])
    batch = torch.tensor([0, 0, 0, 1, 1])
    ei = O.knn_graph(x, 3, batch, loop=False)                  # events smaller than k: all other nodes of the event
    assert ei.tolist() == [[1, 2, 0, 2, 1, 0, 4, 3], [0, 0, 1, 1, 2, 2, 3, 4]]
    assert torch.equal(O.knn_graph(x, 3, batch, loop=False, flow="target_to_source"), ei.flip(0))
    ei = O.knn_graph(x, 2, batch, loop=True)                   # loop=True: self first (distance 0)
    assert ei[0, ei[1] == 1].tolist() == [1, 0]
    assert O.knn_graph(torch.zeros(0, 4), 3, torch.zeros(0, dtype=torch.int64)).shape == (2, 0)
