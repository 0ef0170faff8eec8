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
    assert tr._implicit(training=False) and tr._implicit(training=True)
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


@pytest.mark.parametrize("prec", ["bf16", "tf32"])
@pytest.mark.parametrize("shape,sizes,squeeze,max_nn", [("dytt", [300, 1, 0, 260, 2], 1.0, 255), ("highpu", [1500], 0.2, 255),
                                                        ("dytt", [700], 0.3, 16)])
def test_tc_geo_argmax_variant(prec, shape, sizes, squeeze, max_nn):
    """The arg-max variant (training): same values as the plain kernel bit for bit; every recorded source is an edge
    of the reference graph; and the source attains the maximum of the operand-rounding MODEL of the kernel (fp64
    products of the rounded operands) up to the fp32 accumulation error."""
    b, etaphi = _events(shape, sizes, 23, squeeze)
    ei = O.radius_graph(etaphi, 0.4, b.batch, loop=True, max_num_neighbors=max_nn)
    n = etaphi.shape[0]
    torch.manual_seed(6)
    x = torch.randn(n, 32)
    lin = torch.nn.Sequential(torch.nn.Linear(64, 32))
    geo = ops.build_geo(etaphi.to(DEV), 0.4, b.ptr.to(DEV).int(), loop=True, max_num_neighbors=max_nn)
    perm = geo.perm[:n].long()
    Wd, bd = lin[0].weight.detach().to(DEV), lin[0].bias.detach().to(DEV)
    plain = ops.edgeconv_geo(x.to(DEV)[perm], geo, Wd, bd, aggr="max", precision=prec)
    got, arg = ops.edgeconv_geo(x.to(DEV)[perm], geo, Wd, bd, aggr="max", precision=prec, want_arg=True)
    assert torch.equal(plain, got)
    got, arg, perm = got.cpu(), arg.cpu().long(), perm.cpu()
    src = perm[arg.clamp_min(0)]                       # internal -> input node ids
    tgt = perm[:, None].expand(-1, 32)
    edges = set((ei[0] * n + ei[1]).tolist())
    live = arg >= 0
    assert bool(live.all())                            # loop=True: every target has at least itself
    assert set((src * n + tgt).flatten().tolist()) <= edges
    # model of the kernel in fp64: W1 x_i + b + round(W2) round(x_j - x_i)
    W = lin[0].weight.detach().double()
    w2 = O.round_operand(lin[0].weight.detach()[:, 32:], prec).double()
    xi = x[tgt[:, 0]].double()
    base = xi @ W[:, :32].t() + lin[0].bias.detach().double()
    d = O.round_operand((x[src] - x[tgt]).float(), prec).double()         # [n, 32 (channel), 32 (k)]
    val_at_arg = base + torch.einsum("nck,ck->nc", d, w2)
    scale = got.abs().max().item()
    assert (val_at_arg - got.double()).abs().max().item() < 2e-5 * scale  # the output IS the message of the recorded source
    conv = O.EdgeConv(lin.double(), "max"); conv.operand_rounding = prec
    best = conv(x.double(), ei).detach()[perm]
    assert (best - got.double()).abs().max().item() < 2e-5 * scale        # and that message is the maximum


@pytest.mark.parametrize("prec,gtol", [("tf32", 2e-2), ("bf16", 1.5e-1)])
def test_tc_geo_training_step_on_the_implicit_graph(prec, gtol):
    """train.py:49-52 with the tcgen05 edge contraction on the implicit graph: loss within the forward bound, every
    gradient within a stated looser bound of the fp32 oracle's (arg-max flips of near-tied messages move individual
    gradient entries: the bound is on the tensor's max-abs), BatchNorm running statistics, and run-to-run identity."""
    torch.manual_seed(0)
    b = synth.make_batch(3, "dytt", seed=41, sizes=[300, 250, 280])
    ref = O.Net(8, 3); ours = gnet.Net(8, 3); ours.load_state_dict(ref.state_dict()); ours.to(DEV)
    sd0 = {k: v.clone() for k, v in ref.state_dict().items()}
    x_cont, x_cat, etaphi = O.prepare_inputs(b.x)
    ei_ref = O.radius_graph(etaphi, 0.4, b.batch, loop=True, max_num_neighbors=255)
    ref.train()
    loss_ref = O.loss_fn(ref(x_cont, x_cat, ei_ref, b.batch), b.x, b.y, b.batch)
    loss_ref.backward()
    ours.graphnet.edge_precision = prec
    tr = GraphMETTrainer(ours, lr=0.0, weight_decay=0.0, graph="implicit")
    assert tr._implicit(training=True)
    loss = tr.train_step(b.to(DEV))
    torch.cuda.synchronize()
    assert abs(loss.item() - loss_ref.item()) < gtol * abs(loss_ref.item())
    grads = {k: g.clone() for (k, _), g in zip(ours.graphnet.named_parameters(), tr._gviews)}
    for k, p in ref.graphnet.named_parameters():
        scale = p.grad.abs().max().clamp_min(1e-12)
        if k.endswith("nn.0.bias"):
            scale = dict(ref.graphnet.named_parameters())[k.replace("nn.0.bias", "nn.0.weight")].grad.abs().max()
        err = float((grads[k].cpu() - p.grad).abs().max() / scale)
        assert err < gtol, (k, err)
    for k, v in ours.state_dict().items():
        if "running" in k:
            assert torch.allclose(v.cpu(), ref.state_dict()[k], rtol=5e-2, atol=5e-3), k
    # a second trainer on the same weights: bit-identical gradients
    ours2 = gnet.Net(8, 3); ours2.load_state_dict(sd0); ours2.to(DEV); ours2.graphnet.edge_precision = prec
    tr2 = GraphMETTrainer(ours2, lr=0.0, weight_decay=0.0, graph="implicit")
    tr2.train_step(b.to(DEV))
    assert torch.equal(tr2.flat_grad, tr.flat_grad)
