"""k-NN graph kernel (csrc/knn.cu) against oracle/knn_bruteforce.c — SURVEY.md §8 rows f4 / f1:
``knn_graph(emb, k=20, batch, loop=True)`` (model/graph_met_network.py:63 sketch) and
``knn_graph(x, k, batch, loop=False)`` (model/dynamic_reduction_network.py:86,95).  Index work: bit-exact."""
import pytest
import torch

from deepmetv2_b200 import ops
from deepmetv2_b200.model import graph_met_network as G
from oracle import graphmet_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _batch(sizes):
    return torch.repeat_interleave(torch.arange(len(sizes)), torch.tensor(sizes))


@pytest.mark.parametrize("dim,k,loop", [(2, 4, False), (3, 8, True), (8, 8, False), (11, 16, True), (32, 20, True),
                                        (32, 20, False), (64, 8, False), (64, 32, True), (20, 31, False)])
def test_knn_graph_exact_vs_bruteforce(dim, k, loop):
    sizes = [300, 1, 0, 257, 2, 33, 700, 0, 129]
    batch = _batch(sizes)
    g = torch.Generator().manual_seed(dim * 100 + k)
    x = torch.randn(sum(sizes), dim, generator=g)
    want = O.knn_graph(x, k, batch, loop=loop)
    got = ops.knn_graph(x.to(DEV), k, batch.to(DEV), loop=loop)
    assert got.dtype == torch.int64 and torch.equal(got.cpu(), want)
    t2s = ops.knn_graph(x.to(DEV), k, batch.to(DEV), loop=loop, flow="target_to_source")
    assert torch.equal(t2s.cpu(), want.flip(0))


@pytest.mark.parametrize("loop", [False, True])
def test_knn_graph_ties_and_duplicates(loop):
    """Integer lattice points: most distances tie, many exact duplicates (k+1 kept when self is not among the best)."""
    sizes = [90, 40, 5, 200]
    batch = _batch(sizes)
    g = torch.Generator().manual_seed(11)
    x = torch.randint(-2, 3, (sum(sizes), 3), generator=g).float()
    want = O.knn_graph(x, 6, batch, loop=loop)
    got = ops.knn_graph(x.to(DEV), 6, batch.to(DEV), loop=loop)
    assert torch.equal(got.cpu(), want)
    z = ops.knn_graph(torch.zeros(6, 2, device=DEV), 2, None, loop=False)
    assert torch.bincount(z[1].cpu(), minlength=6).tolist() == [2, 2, 2, 3, 3, 3]


def test_knn_graph_one_large_event_and_strided_rows():
    """A high-pile-up sized event (many candidate tiles per query) read from a wider row (stride > dim)."""
    g = torch.Generator().manual_seed(5)
    wide = torch.randn(5000, 40, generator=g)
    x = wide[:, :32]
    want = O.knn_graph(x.contiguous(), 20, None, loop=True)
    got = ops.knn_graph(wide.to(DEV)[:, :32], 20, None, loop=True)
    assert torch.equal(got.cpu(), want)
    again = ops.knn_graph(wide.to(DEV)[:, :32], 20, None, loop=True)
    assert torch.equal(got, again)


def test_knn_graph_rejects_what_the_kernels_do_not_cover():
    x = torch.randn(10, 65, device=DEV)
    with pytest.raises(NotImplementedError):
        ops.knn_graph(x, 4)
    with pytest.raises(NotImplementedError):
        ops.knn_graph(x[:, :4].contiguous(), 32, loop=False)
    assert ops.knn_graph(torch.zeros(0, 4, device=DEV), 3, torch.zeros(0, dtype=torch.int64, device=DEV)).shape == (2, 0)


@pytest.mark.parametrize("aggr", ["max", "add", "mean"])
def test_edgeconv_on_dynamic_knn_graph_matches_oracle(aggr):
    """graph_met_network.py:63 (sketched): EdgeConv over knn_graph(emb, k=20, loop=True) in the 32-d embedding."""
    sizes = [400, 350, 1, 90]
    batch = _batch(sizes)
    torch.manual_seed(2)
    emb = torch.randn(sum(sizes), 32)
    lin = torch.nn.Sequential(torch.nn.Linear(64, 32))
    ei_ref = O.knn_graph(emb, 20, batch, loop=True)
    want = O.EdgeConv(lin, aggr)(emb, ei_ref).detach()
    ei = ops.knn_graph(emb.to(DEV), 20, batch.to(DEV), loop=True)
    assert torch.equal(ei.cpu(), ei_ref)
    conv = G.EdgeConv(torch.nn.Sequential(torch.nn.Linear(64, 32)), aggr).to(DEV)
    conv.nn.load_state_dict(lin.state_dict())
    with torch.no_grad():
        got = conv(emb.to(DEV), ei)
    err = (got.cpu() - want).abs().max().item() / want.abs().max().item()
    assert err < 1e-5, err


def test_build_knn_is_sync_free_and_feeds_edgeconv():
    """ops.build_knn: CSR at its upper-bound capacity, E left on the device; same rows as knn_graph; EdgeConv takes it."""
    sizes = [300, 5, 0, 260]
    batch = _batch(sizes)
    ptr = torch.tensor([0, 300, 305, 305, 565], dtype=torch.int32, device=DEV)
    torch.manual_seed(4)
    emb = torch.randn(sum(sizes), 32)
    for loop in (True, False):
        want = O.knn_graph(emb, 8, batch, loop=loop)
        g = ops.build_knn(emb.to(DEV), 8, ptr, loop=loop)
        E = int(g.rowptr[-1])
        assert E == want.shape[1] and g.col.numel() >= E
        assert torch.equal(g.col[:E].long().cpu(), want[0])
        deg = (g.rowptr[1:] - g.rowptr[:-1]).long().cpu()
        assert torch.equal(torch.repeat_interleave(torch.arange(len(deg)), deg), want[1])
    lin = torch.nn.Sequential(torch.nn.Linear(64, 32))
    ref = O.EdgeConv(lin, "max")(emb, O.knn_graph(emb, 8, batch, loop=True)).detach()
    conv = G.EdgeConv(torch.nn.Sequential(torch.nn.Linear(64, 32)), "max").to(DEV)
    conv.nn.load_state_dict(lin.state_dict())
    with torch.no_grad():
        got = conv(emb.to(DEV), ops.build_knn(emb.to(DEV), 8, ptr, loop=True))
    assert (got.cpu() - ref).abs().max().item() / ref.abs().max().item() < 1e-5
