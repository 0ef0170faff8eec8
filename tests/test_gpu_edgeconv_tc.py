"""tcgen05 / TMEM EdgeConv variants (tf32, bf16 operands, fp32 accumulate) against the oracle.
Stated tolerances: these are NOT the fp32 parity path (north_star: tensor-core variants are reported
separately under a looser bound)."""
import numpy as np
import pytest
import torch

from deepmetv2_b200 import ops, synth
from deepmetv2_b200.model import net as gnet
from deepmetv2_b200.model.graph_met_network import EdgeConv
from oracle import graphmet_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"
# relative to the output's max-abs: tf32 has 10 mantissa bits (rna), bf16 has 7
TOL = {"tf32": 2e-3, "bf16": 1.5e-2}


def _graph(sizes, seed, squeeze=1.0):
    b = synth.make_batch(len(sizes), "dytt", seed=seed, sizes=sizes)
    x = b.x.clone(); x[:, 3] *= squeeze
    etaphi = O.prepare_inputs(x)[2]
    ei = O.radius_graph(etaphi, 0.4, b.batch, loop=True, max_num_neighbors=255)
    return b, ei


@pytest.mark.parametrize("prec", ["tf32", "bf16"])
@pytest.mark.parametrize("aggr", ["max", "add", "mean"])
@pytest.mark.parametrize("sizes,squeeze", [([200, 180], 1.0), ([1, 3, 700, 0, 40], 1.0), ([900], 0.1)])
def test_tc_edgeconv_matches_oracle(prec, aggr, sizes, squeeze):
    b, ei = _graph(sizes, seed=17, squeeze=squeeze)
    n = b.x.shape[0]
    torch.manual_seed(5)
    x = torch.randn(n, 32)
    lin = torch.nn.Sequential(torch.nn.Linear(64, 32))
    want = O.EdgeConv(lin, aggr)(x, ei).detach()
    conv = EdgeConv(torch.nn.Sequential(torch.nn.Linear(64, 32)), aggr).to(DEV)
    conv.nn.load_state_dict(lin.state_dict())
    conv.precision = prec
    with torch.no_grad():
        got = conv(x.to(DEV), ei.to(DEV)).cpu()
    torch.cuda.synchronize()
    err = (got - want).abs().max().item() / want.abs().max().item()
    assert err < TOL[prec], (prec, aggr, err)


@pytest.mark.parametrize("prec", ["tf32", "bf16"])
def test_tc_isolated_nodes_and_loopless_graph(prec):
    """loop=False leaves isolated nodes with empty rows (-> 0, like PyG)."""
    b = synth.make_batch(2, "tiny", seed=3)
    etaphi = O.prepare_inputs(b.x)[2]
    ei = O.radius_graph(etaphi, 0.4, b.batch, loop=False, max_num_neighbors=255)
    n = b.x.shape[0]
    torch.manual_seed(6)
    x = torch.randn(n, 32)
    lin = torch.nn.Sequential(torch.nn.Linear(64, 32))
    want = O.EdgeConv(lin, "max")(x, ei).detach()
    conv = EdgeConv(torch.nn.Sequential(torch.nn.Linear(64, 32)), "max").to(DEV)
    conv.nn.load_state_dict(lin.state_dict()); conv.precision = prec
    with torch.no_grad():
        got = conv(x.to(DEV), ei.to(DEV)).cpu()
    deg = torch.bincount(ei[1], minlength=n)
    assert bool((got[deg == 0] == 0).all())
    assert (got - want).abs().max().item() / want.abs().max().item() < TOL[prec]


@pytest.mark.parametrize("prec,wtol", [("tf32", 3e-3), ("bf16", 3e-2)])
def test_tc_network_forward_and_training_step(prec, wtol):
    """Whole network with the tcgen05 edge contraction: weights within the stated looser tolerance of
    the fp32 oracle; a training step runs (the backward is precision-independent given arg-max)."""
    torch.manual_seed(0)
    b = synth.make_batch(3, "dytt", seed=41, sizes=[300, 250, 280])
    ref = O.Net(8, 3); ours = gnet.Net(8, 3); ours.load_state_dict(ref.state_dict()); ours.to(DEV)
    ours.graphnet.edge_precision = prec
    g = b.to(DEV)
    x_cont, x_cat, etaphi = O.prepare_inputs(b.x)
    ei_ref = O.radius_graph(etaphi, 0.4, b.batch, loop=True, max_num_neighbors=255)
    ei = ops.radius_graph(etaphi.to(DEV), 0.4, g.batch, loop=True, max_num_neighbors=255)
    ref.eval(); ours.eval()
    with torch.no_grad():
        w_ref = ref(x_cont, x_cat, ei_ref, b.batch)
        w = ours(g.x[:, :8], g.x[:, 8:].long(), ei, g.batch)
    rel = ((w.cpu() - w_ref).abs() / w_ref.abs()).max().item()
    assert rel < wtol, rel
    ours.train(); ref.train()
    w = ours(g.x[:, :8], g.x[:, 8:].long(), ei, g.batch)
    loss = gnet.loss_fn(w, g.x, g.y, g.batch)
    loss.backward()
    w_ref = ref(x_cont, x_cat, ei_ref, b.batch)
    loss_ref = O.loss_fn(w_ref, b.x, b.y, b.batch)
    loss_ref.backward()
    assert abs(loss.item() - loss_ref.item()) < 10 * wtol * abs(loss_ref.item())
    gw = ours.graphnet.output[0].weight.grad.cpu()
    gw_ref = ref.graphnet.output[0].weight.grad
    assert (gw - gw_ref).abs().max() / gw_ref.abs().max() < 20 * wtol
