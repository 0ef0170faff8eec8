"""Property tests pinning the radius-graph restatement."""
import numpy as np
import torch

from deepmetv2_b200 import synth
from oracle import graphmet_oracle as O


def _etaphi(b):
    return O.prepare_inputs(b.x)[2]


def test_c_bruteforce_equals_torch_bruteforce():
    b = synth.make_batch(4, "dytt", seed=3, sizes=[200, 0, 1, 350])
    ep = _etaphi(b)
    a = O.radius_graph(ep, 0.4, b.batch, loop=True, max_num_neighbors=255)
    t = O.radius_graph_torch(ep, 0.4, b.batch, loop=True, max_num_neighbors=255)
    # non-fused vs fused d^2 may differ only for boundary-ambiguous pairs
    if synth.boundary_ambiguous_pairs(ep, b.ptr) == 0:
        assert torch.equal(a, t)
    a2 = O.radius_graph(ep, 0.4, b.batch, loop=False, max_num_neighbors=255)
    assert a2.shape[1] == a.shape[1] - ep.shape[0]
    assert bool((a2[0] != a2[1]).all())


def test_order_is_target_then_source_ascending():
    b = synth.make_batch(3, "dytt", seed=5, sizes=[300, 280, 310])
    ei = O.radius_graph(_etaphi(b), 0.4, b.batch, loop=True, max_num_neighbors=255)
    key = ei[1] * (ei.max() + 1) + ei[0]
    assert bool((key[1:] > key[:-1]).all())
    # edges never cross events
    assert torch.equal(b.batch[ei[0]], b.batch[ei[1]])


def test_cap_keeps_lowest_source_indices():
    n = 400
    pos = torch.zeros(n, 2)
    pos[:, 0] = torch.linspace(0, 0.01, n)  # all within r of each other
    ei = O.radius_graph(pos, 0.4, None, loop=True, max_num_neighbors=255)
    assert ei.shape[1] == n * 255
    for t in (0, 100, 399):
        src = ei[0][ei[1] == t]
        assert torch.equal(src, torch.arange(255))  # self-loop of node 399 is dropped by the cap
    ei2 = O.radius_graph(pos, 0.4, None, loop=False, max_num_neighbors=255)
    src = ei2[0][ei2[1] == 3]
    assert src.numel() == 255 and 3 not in src.tolist() and src.max() == 255


def test_strict_inequality_and_no_phi_wrap():
    pos = torch.tensor([[0.0, 0.0], [0.4, 0.0], [0.39999, 0.0], [0.0, 3.1], [0.0, -3.1]])
    ei = O.radius_graph(pos, 0.4, None, loop=True, max_num_neighbors=255)
    pairs = set(map(tuple, ei.t().tolist()))
    assert (1, 0) not in pairs and (2, 0) in pairs
    assert (3, 4) not in pairs  # train.py:47 "problem ... at the +/- pi boundary": no wrap in parity mode
    eiw = O.radius_graph(pos, 0.4, None, loop=True, max_num_neighbors=255, phi_wrap=True)
    assert (3, 4) in set(map(tuple, eiw.t().tolist()))


def test_scatter_add_semantics():
    src = torch.tensor([1.0, 2.0, 3.0, 4.0])
    idx = torch.tensor([0, 0, 2, 2])
    assert torch.equal(O.scatter_add(src, idx), torch.tensor([3.0, 0.0, 7.0]))
    assert O.scatter_add(src, idx, dim_size=5).numel() == 5
