"""CPU: the DRN oracle (oracle/drn_oracle.py) against the fixtures generated from the reference's own
model/dynamic_reduction_network.py (oracle/make_golden_drn.py) and against hand-computable cases of the upstream
operators it restates (row f1)."""
import os
import types

import numpy as np
import pytest
import torch

from oracle import drn_oracle as D

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def load_case(name):
    z = np.load(os.path.join(GOLDEN, f"golden_drn_{name}.npz"))
    sd = {k[4:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("sd::")}
    return z, sd


@pytest.mark.parametrize("name", ["tiny", "small"])
@pytest.mark.parametrize("mode", ["eval", "train"])
def test_restatement_reproduces_reference_fixture(name, mode):
    torch.set_num_threads(1)
    z, sd = load_case(name)
    m = D.DynamicReductionNetwork(input_dim=11, hidden_dim=64, k=8, output_dim=2, aggr="max",
                                  norm=torch.tensor(D.DRN_NORM))
    m.load_state_dict(sd, strict=True)
    m.train(mode == "train")
    rec = {}
    out = m(types.SimpleNamespace(x=torch.from_numpy(z["x"]), batch=torch.from_numpy(z["batch"])), {"record": rec})
    for k in ("edge_index1", "cluster1", "batch1p", "edge_index2", "cluster2"):
        assert np.array_equal(rec[k].numpy(), z[f"{mode}::{k}"]), k
    for k in ("x0", "x1", "weight1", "x1p", "x2", "weight2"):
        assert np.array_equal(rec[k].numpy(), z[f"{mode}::{k}"]), k
    assert np.array_equal(out.detach().numpy(), z[f"{mode}::out"])
    if mode == "train":
        for k, v in m.state_dict().items():
            if "running" in k or "num_batches" in k:
                assert np.array_equal(v.numpy(), z["sd_after::" + k] if "sd_after::" + k in z.files
                                      else z[f"train::sd_after::{k}"]), k


def test_wrapper_matches_fixture():
    z, sd = load_case("tiny")
    net = D.DRNNet()
    net.drn.load_state_dict(sd)
    net.eval()
    out = net(types.SimpleNamespace(x=torch.from_numpy(z["x"]), batch=torch.from_numpy(z["batch"])))
    assert np.allclose(out.detach().numpy(), z["eval::net_out"], rtol=0, atol=0)


def test_to_undirected_and_normalized_cut_hand_case():
    ei = torch.tensor([[1, 2, 2, 0], [0, 0, 1, 1]])          # 1->0, 2->0, 2->1, 0->1
    u = D.to_undirected(ei, 3)
    assert u.tolist() == [[0, 0, 1, 1, 2, 2], [1, 2, 0, 2, 0, 1]]
    w = D.normalized_cut(u, torch.tensor([1., 2., 1., 3., 2., 3.]), 3)
    assert torch.allclose(w, torch.tensor([1., 2., 1., 3., 2., 3.]))     # every degree is 2: 1/2 + 1/2
    path = D.to_undirected(torch.tensor([[0, 1], [1, 2]]), 3)            # 0-1-2: degrees 1, 2, 1
    wp = D.normalized_cut(path, torch.ones(4), 3)
    assert wp.tolist() == [1.5, 1.5, 1.5, 1.5]


def test_graclus_hand_case_and_validity():
    # path 0-1-2-3 with weights (0,1)=1, (1,2)=5, (2,3)=1: identity order visits 0 first -> {0,1}, then 2 -> {2,3}
    ei = D.to_undirected(torch.tensor([[0, 1, 2], [1, 2, 3]]), 4)
    w = torch.tensor([1., 1., 5., 5., 1., 1.])
    assert D.graclus(ei, w, 4).tolist() == [0, 0, 2, 2]
    # visiting 1 first takes the heavy edge
    assert D.graclus(ei, w, 4, node_perm=[1, 0, 2, 3]).tolist() == [0, 1, 1, 3]
    # ties go to the LATER entry of the row; negative weights never match
    star = D.to_undirected(torch.tensor([[1, 2, 3], [0, 0, 0]]), 4)
    assert D.graclus(star, torch.ones(6), 4).tolist() == [0, 1, 2, 0]
    assert D.graclus(star, -torch.ones(6), 4).tolist() == [0, 1, 2, 3]
    # random graph: clusters have size <= 2, pairs are edges, ids are the pair minimum, matching is maximal
    g = torch.Generator().manual_seed(3)
    ei = D.to_undirected(torch.randint(0, 60, (2, 200), generator=g), 60)
    ei = ei[:, ei[0] != ei[1]]
    w = torch.rand(ei.shape[1], generator=g)
    c = D.graclus(ei, w, 60)
    edges = set(map(tuple, ei.t().tolist()))
    members = {}
    for n, cid in enumerate(c.tolist()):
        members.setdefault(cid, []).append(n)
    for cid, ms in members.items():
        assert len(ms) <= 2 and min(ms) == cid
        if len(ms) == 2:
            assert (ms[0], ms[1]) in edges
    single = {ms[0] for ms in members.values() if len(ms) == 1}
    assert not any(a in single and b in single for a, b in edges)


def test_pooling_hand_case():
    cluster = torch.tensor([0, 0, 2, 3, 3, 5])
    inv, perm = D.consecutive_cluster(cluster)
    assert inv.tolist() == [0, 0, 1, 2, 2, 3] and perm.tolist() == [1, 2, 4, 5]
    x = torch.tensor([[1., -1.], [0., 2.], [3., 3.], [-4., 0.], [-5., 1.], [7., 7.]])
    px, pb = D.max_pool_x(cluster, x, torch.tensor([0, 0, 0, 1, 1, 1]))
    assert px.tolist() == [[1., 2.], [3., 3.], [-4., 1.], [7., 7.]] and pb.tolist() == [0, 0, 1, 1]
    assert D.global_max_pool(px, pb).tolist() == [[3., 3.], [7., 7.]]
