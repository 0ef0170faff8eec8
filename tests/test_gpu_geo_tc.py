"""Per-edge tcgen05 EdgeConv on the IMPLICIT radius graph (csrc/geo_tc.cu: every consumer warp drives its own
TMEM pipeline).  Edge-set exactness is checked bit for bit through integer hashes (small integers survive the
bf16 / tf32 rounding of x_j - x_i exactly); values are checked against the oracle under the stated tensor-core
tolerances (north_star: tf32 / bf16 variants are reported under a looser bound than the fp32 parity path)."""
import pytest
import torch

from deepmetv2_b200 import ops, synth
from deepmetv2_b200.model import net as gnet
from deepmetv2_b200.trainer import GraphMETTrainer
from oracle import graphmet_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"
TOL = {"tf32": 2e-3, "bf16": 1.5e-2}          # relative to the output's max-abs (same bounds as test_gpu_edgeconv_tc.py)
PRIMES = [251, 241, 239, 233, 229, 227, 223, 211, 199, 197, 193, 191, 181, 179, 173, 167, 163, 157, 151, 149, 139,
          137, 131, 127, 113, 109, 107, 103, 101, 97, 89]


def _hash_features(ptr):
    n = int(ptr[-1])
    local = torch.cat([torch.arange(int(b - a)) for a, b in zip(ptr[:-1], ptr[1:])]) if n else torch.zeros(0)
    x = torch.ones(n, 32)
    for c in range(1, 32):
        x[:, c] = (local * (c + 1)) % PRIMES[c - 1]
    return x


def _events(shape, sizes, seed, squeeze=1.0):
    b = synth.make_batch(len(sizes), shape, seed=seed, sizes=sizes)
    x = b.x.clone(); x[:, 3] *= squeeze
    etaphi = O.prepare_inputs(x)[2]
    return b, etaphi


@pytest.mark.parametrize("prec", ["bf16", "tf32"])
@pytest.mark.parametrize("shape,sizes,seed,squeeze,loop,max_nn", [
    ("dytt", [300, 1, 0, 260, 2, 411], 3, 1.0, True, 255),
    ("highpu", [900, 1200], 4, 1.0, True, 255),
    ("highpu", [2500], 5, 0.15, True, 255),          # squeezed in eta: multi-chunk neighbourhoods, degrees over the cap
    ("dytt", [500, 333], 6, 1.0, False, 255),        # loop=False: self excluded, isolated nodes -> 0
    ("dytt", [700], 7, 0.3, True, 16),               # small cap: most targets capped
])
def test_tc_geo_edge_set_is_exact(prec, shape, sizes, seed, squeeze, loop, max_nn):
    """W = [I | I], b = 0, aggr = add: out_i = deg * x_i + sum_j (x_j - x_i) = sum_j x_j, all small integers: exact."""
    b, etaphi = _events(shape, sizes, seed, squeeze)
    ptr = b.ptr.long()
    pos_d, ptr_d = etaphi.to(DEV), b.ptr.to(DEV).int()
    g, ei = ops.build_graph(pos_d, 0.4, ptr_i32=ptr_d, loop=loop, max_num_neighbors=max_nn, want_edge_index=True)
    geo = ops.build_geo(pos_d, 0.4, ptr_d, loop=loop, max_num_neighbors=max_nn)
    N = etaphi.shape[0]
    perm = geo.perm[:N].long()
    x = _hash_features(ptr)
    want = torch.zeros(N, 32, dtype=torch.float64)
    eic = ei.cpu()
    want.index_add_(0, eic[1], x[eic[0]].double())
    W = torch.zeros(32, 64); W[:, :32] = torch.eye(32); W[:, 32:] = torch.eye(32)
    got = ops.edgeconv_geo(x.to(DEV)[perm], geo, W.to(DEV), torch.zeros(32, device=DEV), aggr="add", precision=prec)
    torch.cuda.synchronize()
    assert torch.equal(got.double().cpu(), want[perm.cpu()]), "tcgen05 geo path saw a different edge set"


@pytest.mark.parametrize("prec", ["bf16", "tf32"])
@pytest.mark.parametrize("aggr", ["max", "add", "mean"])
@pytest.mark.parametrize("shape,sizes,squeeze", [("dytt", [200, 180], 1.0), ("dytt", [1, 3, 700, 0, 40], 1.0),
                                                 ("highpu", [1500], 0.2)])
def test_tc_geo_edgeconv_matches_oracle(prec, aggr, shape, sizes, squeeze):
    b, etaphi = _events(shape, sizes, 17, squeeze)
    ei = O.radius_graph(etaphi, 0.4, b.batch, loop=True, max_num_neighbors=255)
    n = etaphi.shape[0]
    torch.manual_seed(5)
    x = torch.randn(n, 32)
    lin = torch.nn.Sequential(torch.nn.Linear(64, 32))
    want = O.EdgeConv(lin, aggr)(x, ei).detach()
    geo = ops.build_geo(etaphi.to(DEV), 0.4, b.ptr.to(DEV).int(), loop=True, max_num_neighbors=255)
    perm = geo.perm[:n].long()
    got = ops.edgeconv_geo(x.to(DEV)[perm], geo, lin[0].weight.detach().to(DEV), lin[0].bias.detach().to(DEV),
                           aggr=aggr, precision=prec)
    torch.cuda.synchronize()
    back = torch.empty_like(got); back[perm] = got
    err = (back.cpu() - want).abs().max().item() / want.abs().max().item()
    assert err < TOL[prec], (prec, aggr, err)
    # and run to run bit-identical (fixed-order segmented scan, no atomics)
    again = ops.edgeconv_geo(x.to(DEV)[perm], geo, lin[0].weight.detach().to(DEV), lin[0].bias.detach().to(DEV),
                             aggr=aggr, precision=prec)
    assert torch.equal(got, again)


@pytest.mark.parametrize("prec,wtol", [("tf32", 3e-3), ("bf16", 3e-2)])
def test_tc_geo_network_eval_step(prec, wtol):
    """evaluate.py:81-100 through the fused eval step with the tcgen05 edge contraction on the implicit graph."""
    torch.manual_seed(0)
    b = synth.make_batch(3, "dytt", seed=41, sizes=[300, 250, 280])
    ref = O.Net(8, 3); ours = gnet.Net(8, 3); ours.load_state_dict(ref.state_dict()); ours.to(DEV)
    x_cont, x_cat, etaphi = O.prepare_inputs(b.x)
    ei_ref = O.radius_graph(etaphi, 0.4, b.batch, loop=True, max_num_neighbors=255)
    ref.eval()
    with torch.no_grad():
        w_ref = ref(x_cont, x_cat, ei_ref, b.batch)
        loss_ref = O.loss_fn(w_ref, b.x, b.y, b.batch)
    ours.graphnet.edge_precision = prec
    tr = GraphMETTrainer(ours, graph="implicit")
    assert tr._implicit(training=False) and not tr._implicit(training=True)
    w, loss, met = tr.eval_step(b.to(DEV))
    torch.cuda.synchronize()
    rel = ((w.cpu() - w_ref).abs() / w_ref.abs()).max().item()
    assert rel < wtol, rel
    assert abs(loss.item() - loss_ref.item()) < 10 * wtol * abs(loss_ref.item())


@pytest.mark.parametrize("prec", ["bf16", "tf32"])
def test_tc_geo_edge_set_with_phi_wrap(prec):
    """The opt-in periodic phi grid through the tcgen05 path: neighbourhoods wrap across +-pi."""
    b, etaphi = _events("dytt", [600, 450], 9)
    ptr = b.ptr.long()
    pos_d, ptr_d = etaphi.to(DEV), b.ptr.to(DEV).int()
    g, ei = ops.build_graph(pos_d, 0.4, ptr_i32=ptr_d, loop=True, max_num_neighbors=255, phi_wrap=True,
                            want_edge_index=True)
    geo = ops.build_geo(pos_d, 0.4, ptr_d, loop=True, max_num_neighbors=255, phi_wrap=True)
    N = etaphi.shape[0]
    perm = geo.perm[:N].long()
    x = _hash_features(ptr)
    want = torch.zeros(N, 32, dtype=torch.float64)
    eic = ei.cpu()
    want.index_add_(0, eic[1], x[eic[0]].double())
    W = torch.zeros(32, 64); W[:, :32] = torch.eye(32); W[:, 32:] = torch.eye(32)
    got = ops.edgeconv_geo(x.to(DEV)[perm], geo, W.to(DEV), torch.zeros(32, device=DEV), aggr="add", precision=prec)
    assert torch.equal(got.double().cpu(), want[perm.cpu()])
