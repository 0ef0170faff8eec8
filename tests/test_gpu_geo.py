"""The implicit radius graph (gm_geo_build + cell-staged EdgeConv) must describe EXACTLY the edge set of the
explicit, bit-exact-to-the-oracle CSR path — including the max_num_neighbors cap (lowest source indices survive),
loop=False, phi wrap, duplicates, empty events — and give the same EdgeConv results.

Edge-set exactness is checked through integer hashes: with W = [I | I], b = 0 the layer computes
out_i = sum_{j in N(i)} x_j, and x_j holds small integers (1, idx mod p, ...), so every sum is exact in fp32 and
two different neighbour sets give different rows with overwhelming probability."""
import numpy as np
import pytest
import torch

from deepmetv2_b200 import ops, synth
from oracle import graphmet_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"
PRIMES = [251, 241, 239, 233, 229, 227, 223, 211, 199, 197, 193, 191, 181, 179, 173, 167, 163, 157, 151, 149, 139,
          137, 131, 127, 113, 109, 107, 103, 101, 97, 89]


def _hash_features(ptr):
    """[N,32] small-integer features: channel 0 = 1 (degree), channel c = (local index * (c+1)) mod prime_c."""
    n = int(ptr[-1])
    local = torch.cat([torch.arange(int(b - a)) for a, b in zip(ptr[:-1], ptr[1:])]) if n else torch.zeros(0)
    x = torch.ones(n, 32)
    for c in range(1, 32):
        x[:, c] = (local * (c + 1)) % PRIMES[c - 1]
    return x


def _identity_edgeconv():
    W = torch.zeros(32, 64)
    W[:, :32] = torch.eye(32)      # W1 = I
    W[:, 32:] = torch.eye(32)      # W2 = I  ->  Y = x, P = 0
    return W.to(DEV), torch.zeros(32, device=DEV)


def _check_edge_set(pos, ptr, r, loop, max_nn, phi_wrap=False):
    pos_d, ptr_d = pos.to(DEV), ptr.to(DEV).int()
    g, ei = ops.build_graph(pos_d, r, ptr_i32=ptr_d, loop=loop, max_num_neighbors=max_nn, phi_wrap=phi_wrap,
                            want_edge_index=True)
    geo = ops.build_geo(pos_d, r, ptr_d, loop=loop, max_num_neighbors=max_nn, phi_wrap=phi_wrap)
    N = pos.shape[0]
    perm = geo.perm[:N].long()
    assert torch.equal(torch.sort(perm).values, torch.arange(N, device=DEV)), "node_perm is not a permutation"
    x = _hash_features(ptr)
    want = torch.zeros(N, 32, dtype=torch.float64)
    eic = ei.cpu()
    want.index_add_(0, eic[1], x[eic[0]].double())
    W, b = _identity_edgeconv()
    got = ops.edgeconv_geo(x.to(DEV)[perm], geo, W, b, aggr="add")
    assert torch.equal(got.double().cpu(), want[perm.cpu()]), "implicit graph differs from the explicit edge list"
    stats = geo.stats.tolist()
    assert stats[2] == g.num_capped
    return g, geo


def test_implicit_graph_equals_explicit_edge_list_on_synthetic_events():
    for shape, sizes, seed in [("dytt", [300, 1, 0, 260, 2, 411], 3), ("highpu", [900, 1200], 4)]:
        b = synth.make_batch(len(sizes), shape, seed=seed, sizes=sizes)
        _, _, etaphi = O.prepare_inputs(b.x)
        for loop in (True, False):
            _check_edge_set(etaphi, b.ptr, 0.4, loop, 255)
            _check_edge_set(etaphi, b.ptr, 0.4, loop, 255, phi_wrap=True)


def test_implicit_graph_per_event_binning_path():
    """>= 128 events of ordinary size take the one-CTA-per-event binning kernel (k_bin_event) instead of the four
    grid-wide launches: same edge set, with empty / one-node events, duplicates, a binding cap and phi wrap."""
    gen = torch.Generator().manual_seed(21)
    sizes = [int(v) for v in torch.randint(30, 220, (150,), generator=gen)]
    sizes[3], sizes[77], sizes[149] = 0, 1, 2
    b = synth.make_batch(len(sizes), "dytt", seed=9, sizes=sizes)
    _, _, etaphi = O.prepare_inputs(b.x)
    etaphi[40:60] = etaphi[40:41]                                      # exact duplicates inside event 0
    for loop in (True, False):
        _check_edge_set(etaphi, b.ptr, 0.4, loop, 255)
        g, _ = _check_edge_set(etaphi, b.ptr, 0.8, loop, 8, phi_wrap=True)
        assert g.num_capped > 0


def test_implicit_graph_cap_duplicates_and_dense_cells():
    gen = torch.Generator().manual_seed(11)
    # two dense blobs (degrees far above the cap), duplicates, and a sparse halo, in 3 events
    blob = torch.randn(700, 2, generator=gen) * 0.12
    halo = (torch.rand(300, 2, generator=gen) - 0.5) * torch.tensor([6.0, 6.0])
    ev0 = torch.cat([blob, halo])[torch.randperm(1000, generator=gen)]
    ev1 = torch.cat([blob[:200], blob[:200], halo[:50]])              # exact duplicates
    ev2 = torch.randn(600, 2, generator=gen) * 0.05 + torch.tensor([1.0, 3.1])   # everything in ~one cell, near +pi
    pos = torch.cat([ev0, ev1, ev2]).float().contiguous()
    ptr = torch.tensor([0, 1000, 1450, 2050])
    for max_nn in (255, 16, 1):
        for loop in (True, False):
            g, _ = _check_edge_set(pos, ptr, 0.4, loop, max_nn)
            assert g.num_capped > 0
    _check_edge_set(pos, ptr, 0.4, True, 32, phi_wrap=True)
    _check_edge_set(pos, ptr, 0.05, True, 8)                           # many cells (grid saturates at 32 x 32)


def test_implicit_edgeconv_max_is_bit_identical_to_csr_edgeconv():
    torch.manual_seed(5)
    b = synth.make_batch(3, "highpu", seed=21, sizes=[700, 900, 500])
    _, _, etaphi = O.prepare_inputs(b.x)
    pos_d, ptr_d = etaphi.to(DEV), b.ptr.to(DEV).int()
    g = ops.build_graph(pos_d, 0.4, ptr_i32=ptr_d, loop=True, max_num_neighbors=64)
    geo = ops.build_geo(pos_d, 0.4, ptr_d, loop=True, max_num_neighbors=64)
    N = etaphi.shape[0]
    perm = geo.perm[:N].long()
    x = torch.randn(N, 32, device=DEV)
    lin = torch.nn.Linear(64, 32).to(DEV)
    from deepmetv2_b200.model.graph_met_network import EdgeConv
    for aggr in ("max", "add", "mean"):
        conv = EdgeConv(torch.nn.Sequential(lin), aggr=aggr)
        with torch.no_grad():
            want = conv(x, g)
        got, arg = ops.edgeconv_geo(x[perm], geo, lin.weight.detach(), lin.bias.detach(), aggr=aggr, want_arg=True)
        if aggr == "max":
            assert torch.equal(got, want[perm])                       # max is order-independent: same bits
            # the recorded arg-max source really attains the maximum
            Y = x[perm] @ lin.weight.detach()[:, 32:].T
            P = got - torch.gather(Y, 0, arg.long().clamp_min(0))
            assert torch.allclose(P, (x[perm] @ (lin.weight[:, :32] - lin.weight[:, 32:]).T + lin.bias).detach(),
                                  rtol=1e-4, atol=1e-4)
        else:
            assert torch.allclose(got, want[perm], rtol=1e-5, atol=1e-4)


def test_implicit_path_is_bit_deterministic():
    torch.manual_seed(6)
    from deepmetv2_b200.model import net as gnet
    from deepmetv2_b200.trainer import GraphMETTrainer
    g = synth.make_batch(4, "dytt", seed=61, sizes=[400, 380, 5, 420]).to(DEV)
    losses, grads = [], []
    for _ in range(3):
        torch.manual_seed(7)
        m = gnet.Net(8, 3).to(DEV)
        tr = GraphMETTrainer(m, graph="implicit")
        losses.append(tr.train_step(g).item())
        grads.append(tr.flat_grad.clone())
    assert losses[0] == losses[1] == losses[2]
    assert torch.equal(grads[0], grads[1]) and torch.equal(grads[0], grads[2])
