"""GPU parity of the DynamicReductionNetwork variant (row f1) against oracle/drn_oracle.py and the fixtures generated
from the reference's own model/dynamic_reduction_network.py.

Discrete stages (k-NN graph, to_undirected, graclus, pooling indices) are compared BIT-EXACTLY with the oracle stage
run on the same (GPU-produced) float inputs; float stages are compared on identical graphs / clusters.  Tolerances:
node / edge features <= 1e-4 of the tensor's max-abs (fp32, different summation order), normalized-cut weights
<= 1e-5 relative, outputs <= 1e-4 relative to max-abs."""
import os
import types

import numpy as np
import pytest
import torch

from deepmetv2_b200 import ops, synth
from deepmetv2_b200.model import dynamic_reduction_network as drn
from oracle import drn_oracle as D
from oracle import graphmet_oracle as O

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
DEV = "cuda:0"


def close(a, b, tol, what):
    a, b = a.detach().cpu().double(), b.detach().cpu().double()
    assert a.shape == b.shape, (what, a.shape, b.shape)
    scale = max(float(b.abs().max()), 1e-30) if b.numel() else 1.0
    err = float((a - b).abs().max()) / scale if b.numel() else 0.0
    assert err <= tol, (what, err)
    return err


def rand_events(sizes, dim, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(sum(sizes), dim, generator=g)
    ptr = torch.tensor(np.concatenate([[0], np.cumsum(sizes)]), dtype=torch.int32)
    batch = torch.repeat_interleave(torch.arange(len(sizes)), torch.tensor(sizes))
    return x, ptr, batch


@pytest.mark.parametrize("sizes,k", [([40, 1, 0, 25, 2], 8), ([300, 420, 260], 8), ([7, 5], 16), ([1500], 4)])
def test_to_undirected_is_bit_exact(sizes, k):
    x, ptr, batch = rand_events(sizes, 64, 3)
    x[5] = x[3]                                    # exact duplicates
    ei = ops.knn_graph(x.to(DEV), k, batch.to(DEV), loop=False)
    ref = D.to_undirected(ei.cpu(), x.shape[0])
    got = ops.to_undirected(ei, x.shape[0])
    assert torch.equal(got.cpu(), ref)
    # a directed list with repeated entries and self loops (generic CSR input)
    g = torch.Generator().manual_seed(9)
    raw = torch.randint(0, 50, (2, 400), generator=g)
    assert torch.equal(ops.to_undirected(raw.to(DEV), 50).cpu(), D.to_undirected(raw, 50))


def test_normalized_cut_graclus_and_pool_match_oracle():
    sizes = [200, 1, 150, 333]
    x, ptr, batch = rand_events(sizes, 64, 5)
    N = x.shape[0]
    ei = ops.to_undirected(ops.knn_graph(x.to(DEV), 8, batch.to(DEV), loop=False), N)
    w = ops.normalized_cut_2d(ei, x.to(DEV))
    w_ref = D.normalized_cut_2d(ei.cpu(), x)
    assert float(((w.cpu() - w_ref).abs() / w_ref.abs().clamp_min(1e-30)).max()) <= 1e-5
    # graclus: same weights in, identical clusters out; identity order and an explicit visiting order
    cl, partner = ops.graclus(ei, w, N, ptr_i32=ptr.to(DEV), return_partner=True)
    assert torch.equal(cl.cpu(), D.graclus(ei.cpu(), w.cpu(), N))
    g = torch.Generator().manual_seed(1)
    perm = torch.cat([int(ptr[e]) + torch.randperm(sizes[e], generator=g) for e in range(len(sizes))])
    cl_p = ops.graclus(ei, w, N, ptr_i32=ptr.to(DEV), node_perm=perm.to(DEV))
    assert torch.equal(cl_p.cpu(), D.graclus(ei.cpu(), w.cpu(), N, node_perm=perm.tolist()))
    # ties (integer weights: many equal) and the unweighted variant
    wi = torch.randint(0, 3, (ei.shape[1],), generator=g).float()
    assert torch.equal(ops.graclus(ei, wi.to(DEV), N, ptr_i32=ptr.to(DEV)).cpu(), D.graclus(ei.cpu(), wi, N))
    assert torch.equal(ops.graclus(ei, None, N, ptr_i32=ptr.to(DEV)).cpu(), D.graclus(ei.cpu(), None, N))
    # one warp for the whole batch (no ptr) gives the same matching
    assert torch.equal(ops.graclus(ei, w, N).cpu(), cl.cpu())
    # pooling
    xp, ptr_p, newid = ops.cluster_max_pool(x.to(DEV), cl.to(DEV), partner, ptr.to(DEV))
    xr, br = D.max_pool_x(cl.cpu(), x, batch)
    assert torch.equal(xp.cpu(), xr)
    assert torch.equal(ptr_p.cpu().long(), O.batch_to_ptr(br, len(sizes)).long())
    assert torch.equal(newid.cpu().long(), D.consecutive_cluster(cl.cpu())[0])


def load_case(name):
    z = np.load(os.path.join(GOLDEN, f"golden_drn_{name}.npz"))
    sd = {k[4:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("sd::")}
    return z, sd


def csr_to_ei(rowptr, col):
    n = rowptr.numel() - 1
    row = torch.repeat_interleave(torch.arange(n), (rowptr[1:] - rowptr[:-1]).long())
    return torch.stack([row, col.long()], dim=0)


def stagewise(ours, ref, x, batch, mode):
    """Runs our forward, then checks every stage against the ORACLE stage evaluated on our stage's inputs."""
    ref.train(mode == "train")
    ours.train(mode == "train")
    data = types.SimpleNamespace(x=x.to(DEV), batch=batch.to(DEV))
    out = ours(data)
    it = {k: v.cpu() for k, v in ours.intermediates().items()}
    N, k = x.shape[0], ref.k
    with torch.no_grad():
        close(it["x0"], ref.inputnet(ref.datanorm * x), 1e-5, "x0")
        ei1 = csr_to_ei(it["rowptr1"], it["col1"])
        assert torch.equal(ei1, D.to_undirected(O.knn_graph(it["x0"], k, batch, loop=False), N)), "graph 1"
        assert torch.equal(it["row1"].long(), ei1[0])
        close(it["x1"], ref.edgeconv1(it["x0"], ei1), 1e-4, "x1")
        w1 = D.normalized_cut_2d(ei1, it["x1"])
        assert float(((it["weight1"] - w1).abs() / w1.abs().clamp_min(1e-30)).max()) <= 1e-5
        cl1 = D.graclus(ei1, it["weight1"], N)
        assert torch.equal(it["cluster1"].long(), cl1), "cluster 1"
        xp, bp = D.max_pool_x(cl1, it["x1"], batch)
        assert torch.equal(it["x1p"], xp), "x1p"
        B = int(batch.max()) + 1
        assert torch.equal(it["ptr1p"].long(), O.batch_to_ptr(bp, B).long())
        M = xp.shape[0]
        ei2 = csr_to_ei(it["rowptr2"], it["col2"])
        assert torch.equal(ei2, D.to_undirected(O.knn_graph(xp, k, bp, loop=False), M)), "graph 2"
        close(it["x2"], ref.edgeconv2(xp, ei2), 1e-4, "x2")
        pooled = D.global_max_pool(it["x2"], bp)
        assert torch.equal(it["pooled"], pooled), "pooled"
        close(out, ref.output(pooled).squeeze(-1), 1e-5, "out")
    return out.cpu(), it


@pytest.mark.parametrize("name", ["tiny", "small"])
@pytest.mark.parametrize("mode", ["eval", "train"])
def test_drn_forward_stagewise_and_against_reference_fixture(name, mode):
    z, sd = load_case(name)
    ref = D.DynamicReductionNetwork(input_dim=11, hidden_dim=64, k=8, output_dim=2, aggr="max", norm=torch.tensor(D.DRN_NORM))
    ref.load_state_dict(sd)
    ours = drn.DynamicReductionNetwork(input_dim=11, hidden_dim=64, k=8, output_dim=2, aggr="max",
                                       norm=torch.tensor(D.DRN_NORM))
    ours.load_state_dict(sd, strict=True)
    ours.to(DEV)
    x, batch = torch.from_numpy(z["x"]), torch.from_numpy(z["batch"])
    out, it = stagewise(ours, ref, x, batch, mode)
    # end to end against the fixture of the reference's own file: the discrete stages must coincide (no near-tie flips
    # on these seeds), then the outputs agree to fp32 accuracy
    ei1 = csr_to_ei(it["rowptr1"], it["col1"])
    assert np.array_equal(ei1.numpy(), z[f"{mode}::edge_index1"])
    assert np.array_equal(it["cluster1"].numpy(), z[f"{mode}::cluster1"])
    assert np.array_equal(csr_to_ei(it["rowptr2"], it["col2"]).numpy(), z[f"{mode}::edge_index2"])
    close(it["x1"], torch.from_numpy(z[f"{mode}::x1"]), 1e-4, "x1 vs fixture")
    close(it["x2"], torch.from_numpy(z[f"{mode}::x2"]), 1e-4, "x2 vs fixture")
    close(out, torch.from_numpy(z[f"{mode}::out"]), 1e-4, "out vs fixture")
    if mode == "train":     # BatchNorm over the edge rows: running statistics after one step
        for ec in ("edgeconv1", "edgeconv2"):
            bn = getattr(ours, ec).nn[4]
            close(bn.running_mean, torch.from_numpy(z[f"train::sd_after::{ec}.nn.4.running_mean"]), 1e-5, ec + " mean")
            close(bn.running_var, torch.from_numpy(z[f"train::sd_after::{ec}.nn.4.running_var"]), 1e-4, ec + " var")
            assert int(bn.num_batches_tracked) == int(z[f"train::sd_after::{ec}.nn.4.num_batches_tracked"])


def test_drn_wrapper_and_c4_sized_batch_properties():
    torch.manual_seed(3)
    ref = D.DRNNet()
    ours = drn.DRNNet()
    ours.load_state_dict(ref.state_dict(), strict=True)
    ours.to(DEV).eval()
    ref.eval()
    b = synth.make_batch(32, "znunu", seed=41)
    g = b.to(DEV)
    out = ours(g)
    out2 = ours(g)
    assert torch.equal(out, out2), "run-to-run determinism"
    assert out.shape == (32, 2) and bool((out[:, 0] > 0).all()) and bool((out[:, 1].abs() <= np.pi).all())
    it = ours.drn.intermediates()
    # the undirected graph is symmetric, sorted, free of self loops
    ei = csr_to_ei(it["rowptr1"].cpu(), it["col1"].cpu())
    key = ei[0] * g.x.shape[0] + ei[1]
    assert bool((key[1:] > key[:-1]).all()) and bool((ei[0] != ei[1]).all())
    assert torch.equal(torch.sort(ei[1] * g.x.shape[0] + ei[0]).values, key)
    # event permutation (eval mode): outputs permute with the events
    order = torch.randperm(32, generator=torch.Generator().manual_seed(0)).tolist()
    xs = [b.x[int(b.ptr[e]):int(b.ptr[e + 1])] for e in order]
    pb = synth.EventBatch(torch.cat(xs), b.y[order], torch.repeat_interleave(torch.arange(32), torch.tensor([len(a) for a in xs])),
                          torch.tensor(np.concatenate([[0], np.cumsum([len(a) for a in xs])]), dtype=torch.int32))
    outp = ours(pb.to(DEV))
    assert torch.equal(outp, out[order])
    # four whole events against the oracle (stage-wise inside)
    small = synth.make_batch(4, "znunu", seed=42)
    o, _ = stagewise(ours.drn, ref.drn, small.x, small.batch, "eval")
    with torch.no_grad():
        full = ref(types.SimpleNamespace(x=small.x, batch=small.batch))
    close(ours(small.to(DEV)), full, 1e-3, "wrapper vs oracle end to end")


def _edgeconv_bf16_model(ec, x, ei):
    """Arithmetic model of k_drn_edge_tc in fp64: h1 = ELU(Linear1(cat[x_i, x_j - x_i])) in fp32, rounded to bf16;
    W2 rounded to bf16; exact products, fp32-equivalent accumulation replaced by fp64."""
    import torch.nn.functional as F
    j, i = ei[0], ei[1]
    lin1, lin2, bn = ec.nn[0], ec.nn[2], ec.nn[4]
    h1 = F.elu(lin1(torch.cat([x[i], x[j] - x[i]], 1)))
    h2 = F.elu(O.round_operand(h1, "bf16").double() @ O.round_operand(lin2.weight, "bf16").double().t() + lin2.bias.double())
    msg = bn(h2.float())
    out = msg.new_zeros(x.shape[0], msg.shape[1])
    return out.scatter_reduce(0, i.view(-1, 1).expand(-1, msg.shape[1]), msg, "amax", include_self=False)


@pytest.mark.parametrize("mode", ["eval", "train"])
def test_drn_bf16_tcgen05_edge_mlp(mode):
    """edge_precision = "bf16": the 96 -> 64 per-edge contraction on tcgen05.  Reported under a looser, stated bound:
    EdgeConv outputs within 5e-3 of max-abs of the operand-rounding model (fp32 re-association before the bf16 rounding
    flips individual operands by one bf16 ulp; train-mode BatchNorm divides by the batch std) and within 2e-2 (eval) /
    5e-2 (train) of the fp32 oracle; discrete stages stay exact GIVEN the kernel's own float outputs."""
    z, sd = load_case("small")
    ref = D.DynamicReductionNetwork(input_dim=11, hidden_dim=64, k=8, output_dim=2, aggr="max", norm=torch.tensor(D.DRN_NORM))
    ref.load_state_dict(sd)
    ours = drn.DynamicReductionNetwork(input_dim=11, hidden_dim=64, k=8, output_dim=2, aggr="max", norm=torch.tensor(D.DRN_NORM))
    ours.load_state_dict(sd, strict=True)
    ours.to(DEV)
    ours.edge_precision = "bf16"
    ref.train(mode == "train"); ours.train(mode == "train")
    x, batch = torch.from_numpy(z["x"]), torch.from_numpy(z["batch"])
    with torch.no_grad():        # the tcgen05 variant is forward-only (the backward differentiates the fp32 forward)
        out = ours(types.SimpleNamespace(x=x.to(DEV), batch=batch.to(DEV)))
        out_again = ours(types.SimpleNamespace(x=x.to(DEV), batch=batch.to(DEV))) if mode == "eval" else out
    assert torch.equal(out, out_again)
    it = {k: v.cpu() for k, v in ours.intermediates().items()}
    N = x.shape[0]
    with torch.no_grad():
        ei1 = csr_to_ei(it["rowptr1"], it["col1"])
        assert torch.equal(ei1, D.to_undirected(O.knn_graph(it["x0"], 8, batch, loop=False), N))
        e_model = close(it["x1"], _edgeconv_bf16_model(ref.edgeconv1, it["x0"], ei1), 5e-3, "x1 vs bf16 model")
        ref.train(mode == "train")
        # eval: BatchNorm uses the running statistics; train: it divides by the batch std of h2, which scales the bf16 error
        e_fp32 = close(it["x1"], ref.edgeconv1(it["x0"], ei1), 2e-2 if mode == "eval" else 5e-2, "x1 vs fp32 oracle")
        cl1 = D.graclus(ei1, it["weight1"], N)
        assert torch.equal(it["cluster1"].long(), cl1)
        xp, bp = D.max_pool_x(cl1, it["x1"], batch)
        assert torch.equal(it["x1p"], xp)
        ei2 = csr_to_ei(it["rowptr2"], it["col2"])
        assert torch.equal(ei2, D.to_undirected(O.knn_graph(xp, 8, bp, loop=False), xp.shape[0]))
        close(it["x2"], _edgeconv_bf16_model(ref.edgeconv2, xp, ei2), 5e-3, "x2 vs bf16 model")
        close(out, ref.output(D.global_max_pool(it["x2"], bp)).squeeze(-1), 1e-5, "out")
    print(f"\n[DRN bf16 tcgen05, {mode}] x1 vs operand-rounding model {e_model:.2e} | vs fp32 oracle {e_fp32:.2e}")


def _drn_case(name):
    """(state_dict, x, batch, datanorm): the fixtures of the reference's own file, or a 980-node batch with a datanorm
    that keeps the continuous features visible."""
    if name in ("tiny", "small"):
        z, sd = load_case(name)
        return sd, torch.from_numpy(z["x"]), torch.from_numpy(z["batch"]), torch.tensor(D.DRN_NORM)
    torch.manual_seed(77)
    norm = torch.tensor([1 / 3., 1 / 3., 1 / 3., 1 / 2., 2., 1., 3., 1 / 211., 1., 1 / 7., 1.])
    m = D.DynamicReductionNetwork(input_dim=11, hidden_dim=64, k=8, output_dim=2, aggr="max", norm=norm)
    with torch.no_grad():
        for ec in (m.edgeconv1, m.edgeconv2):
            bn = ec.nn[4]
            bn.weight.uniform_(-1.5, 1.5); bn.bias.uniform_(-0.5, 0.5)
    b = synth.make_batch(3, "dytt", seed=91, sizes=[300, 420, 260])
    return {k: v.clone() for k, v in m.state_dict().items()}, b.x, b.batch, norm


def _edgeconv_selected(ec, x, ei, arg):
    """The oracle's EdgeConv with the max taken at GIVEN edges: out[i, c] = msg[edge(i <- arg[i, c]), c].  Which of two
    (nearly) equal messages is the maximum is an integer decision of the forward, like the graph itself; the oracle
    differentiates through the decisions our forward took."""
    j, i = ei[0], ei[1]
    n = x.shape[0]
    msg = ec.nn(torch.cat([x[i], x[j] - x[i]], 1))
    key = i * n + j
    order = torch.argsort(key)
    deg = torch.bincount(i, minlength=n)
    q = torch.arange(n)[:, None] * n + arg.long().clamp_min(0)
    pos = torch.searchsorted(key[order], q.reshape(-1)).clamp_max(key.numel() - 1)
    eidx = order[pos].view(n, 64)
    assert bool(((key[eidx] == q) | (deg == 0)[:, None]).all()), "a selected source is not an edge of the graph"
    out = msg.gather(0, eidx)
    return torch.where((deg > 0)[:, None], out, torch.zeros_like(out))


@pytest.mark.parametrize("name", ["tiny", "small", "synthetic"])
def test_drn_backward_matches_oracle_autograd(name):
    """gm_drn_backward: all 31 gradients (datanorm, inputnet, both edge MLPs with their BatchNorm-over-edges, output MLP)
    against torch autograd through the oracle, run on the graphs, clusters, pooling winners and arg-max edges our forward
    used (integer decisions of the forward are constants of the differentiation, in the reference too; with the wrapper's
    normalisation -- px / 2950 ... -- candidates that share (pdgId, charge, fromPV) have latent rows that agree to ~1e-4,
    and a maximum between two of them is now and then decided below the fp32 rounding difference of two correct
    forwards: the oracle therefore takes the decisions from our forward and is checked to reproduce its values).
    Bound: 1e-3 of each tensor's max-abs gradient."""
    sd, x, batch, norm = _drn_case(name)
    ref = D.DynamicReductionNetwork(input_dim=11, hidden_dim=64, k=8, output_dim=2, aggr="max", norm=norm.clone())
    ref.load_state_dict(sd)
    ours = drn.DynamicReductionNetwork(input_dim=11, hidden_dim=64, k=8, output_dim=2, aggr="max", norm=norm.clone())
    ours.load_state_dict(sd, strict=True)
    ours.to(DEV)
    ref.train(); ours.train()
    B = int(batch.max()) + 1
    R = torch.randn(B, 2, generator=torch.Generator().manual_seed(7))
    out = ours(types.SimpleNamespace(x=x.to(DEV), batch=batch.to(DEV)))
    (out * R.to(DEV)).sum().backward()
    torch.cuda.synchronize()
    it = {k: v.cpu() for k, v in ours.intermediates().items()}
    ei1, ei2 = csr_to_ei(it["rowptr1"], it["col1"]), csr_to_ei(it["rowptr2"], it["col2"])
    # The oracle, stage by stage on our graphs / clusters.  max_pool picks, per (cluster, channel), the larger of two
    # members: where the two values agree to the last few ulps (both edge maxima sit on the saturated branch of the
    # ELU) the CPU and the GPU forward can order them differently, so the winner -- like the graph and the clusters, an
    # integer decision of the forward -- is taken from OUR forward and the oracle differentiates through that gather.
    arg1, arg2 = it["asel1"], it["asel2"]
    x0 = ref.inputnet(ref.datanorm * x)
    x1 = _edgeconv_selected(ref.edgeconv1, x0, ei1, arg1)
    close(x1, ref.edgeconv1(x0, ei1), 1e-5, "selected edges attain the maximum")
    cl = it["cluster1"].long()
    newid, _ = D.consecutive_cluster(cl)
    M = int(newid.max()) + 1
    lo = torch.full((M,), x.shape[0], dtype=torch.long).scatter_reduce(0, newid, torch.arange(x.shape[0]), "amin")
    hi = torch.zeros(M, dtype=torch.long).scatter_reduce(0, newid, torch.arange(x.shape[0]), "amax")
    ours_x1 = it["x1"]
    win = torch.where(ours_x1[lo] >= ours_x1[hi], lo[:, None].expand(-1, 64), hi[:, None].expand(-1, 64))
    xp = x1.gather(0, win)
    bp = batch[lo]
    assert torch.equal(ours_x1.gather(0, win), it["x1p"])
    x2 = _edgeconv_selected(ref.edgeconv2, xp, ei2, arg2)
    close(x2, ref.edgeconv2(xp, ei2), 1e-5, "selected edges attain the maximum")
    out_ref = ref.output(D.global_max_pool(x2, bp)).squeeze(-1)
    close(out, out_ref, 1e-4, "train-mode output")
    (out_ref * R).sum().backward()
    named = dict(ours.named_parameters())
    worst = ("", 0.0)
    for k, p in ref.named_parameters():
        g, gr = named[k].grad, p.grad
        assert g is not None, k
        err = close(g, gr, 1e-3, "grad " + k)
        worst = max(worst, (k, err), key=lambda t: t[1])
    print(f"\n[DRN backward, {name}] worst gradient error {worst[1]:.2e} of max-abs ({worst[0]})")
    # a second backward on the same inputs: bit-identical gradients (no float atomics anywhere)
    g1 = {k: v.grad.clone() for k, v in named.items()}
    ours.zero_grad()
    sd_bn = {k: v.clone() for k, v in ours.state_dict().items()}
    out2 = ours(types.SimpleNamespace(x=x.to(DEV), batch=batch.to(DEV)))
    (out2 * R.to(DEV)).sum().backward()
    assert all(torch.equal(named[k].grad, g1[k]) for k in g1)


def test_drn_trainer_step_matches_torch_adamw_on_oracle_gradients():
    """DRNTrainer.train_step = forward + fused wrapper / MET loss + backward + flat AdamW: the loss equals the wrapper's
    formula on the oracle side, and after one step every parameter equals torch.optim.AdamW applied to the gradients of
    that loss taken through our own autograd node (itself checked against the oracle above)."""
    import math
    from deepmetv2_b200.trainer import DRNTrainer
    torch.manual_seed(5)
    b = synth.make_batch(4, "znunu", seed=93, sizes=[220, 180, 1, 260])
    net_a, net_b = drn.DRNNet(), drn.DRNNet()
    net_b.load_state_dict(net_a.state_dict())
    net_a.to(DEV); net_b.to(DEV)
    g = b.to(DEV)
    # reference route: our autograd node + torch ops for the wrapper and the loss + torch AdamW
    net_b.train()
    o = net_b(g)                                             # (met, phi)
    px, py = o[:, 0] * torch.cos(o[:, 1]), o[:, 0] * torch.sin(o[:, 1])
    loss_b = 0.5 * ((px - g.y[:, 0]) ** 2 + (py - g.y[:, 1]) ** 2).mean()
    loss_b.backward()
    # fused route
    before = {k: v.detach().clone() for k, v in net_a.named_parameters()}
    tr = DRNTrainer(net_a, lr=1e-3)
    loss_a = tr.train_step(g)
    torch.cuda.synchronize()
    assert abs(loss_a.item() - loss_b.item()) <= 1e-5 * abs(loss_b.item())
    for k, pb in net_b.drn.named_parameters():
        close(tr._gviews[k], pb.grad, 1e-4, "fused-loss gradient " + k)
    # the update: torch.optim.AdamW fed with the fused step's own gradient (AdamW's first step is lr * sign(g): comparing
    # two routes' updates directly would amplify rounding noise of near-zero gradient entries to +-lr)
    clones = {k: torch.nn.Parameter(v.clone()) for k, v in before.items()}
    for k, q in clones.items():
        q.grad = tr._gviews[k.replace("drn.", "", 1)].clone()
    torch.optim.AdamW(list(clones.values()), lr=1e-3).step()
    for k, pa in net_a.named_parameters():
        assert torch.allclose(pa, clones[k], rtol=1e-6, atol=1e-7), k
    for (k, ba), (_, bb) in zip(net_a.named_buffers(), net_b.named_buffers()):
        assert torch.allclose(ba.float(), bb.float(), rtol=1e-6, atol=1e-7), k
    assert math.isfinite(tr.train_step_from_host(b.pin_memory()))


def test_drn_trainer_bf16_forward_straight_through():
    """DRNTrainer with edge_precision = "bf16": tcgen05 forward, fp32 backward routed through the sources the bf16
    forward selected.  Reported under a looser bound: loss within 2e-2 of the all-fp32 step, the flat gradient finite,
    run-to-run identical and aligned with the fp32 gradient (cosine > 0.9)."""
    from deepmetv2_b200.trainer import DRNTrainer
    torch.manual_seed(6)
    b = synth.make_batch(4, "znunu", seed=94, sizes=[400, 380, 500, 450]).to(DEV)
    nets = [drn.DRNNet() for _ in range(3)]
    for n in nets[1:]:
        n.load_state_dict(nets[0].state_dict())
    for n in nets:
        n.to(DEV)
    nets[1].drn.edge_precision = nets[2].drn.edge_precision = "bf16"
    trs = [DRNTrainer(n, lr=0.0, weight_decay=0.0) for n in nets]
    losses = [t.train_step(b) for t in trs]
    torch.cuda.synchronize()
    assert abs(losses[1].item() - losses[0].item()) <= 2e-2 * abs(losses[0].item())
    assert torch.equal(trs[1].flat_grad, trs[2].flat_grad)
    # the two forwards take different integer decisions downstream of the rounded activations (graclus pairs, the stage-2
    # k-NN graph, arg-max edges), so single tensors are not comparable entry by entry; the step as a whole points the
    # same way
    g32, g16 = trs[0].flat_grad, trs[1].flat_grad
    assert torch.isfinite(g16).all()
    cos = float(torch.dot(g32, g16) / (g32.norm() * g16.norm()))
    print(f"\n[DRN bf16-forward training] loss {losses[1].item():.5f} vs fp32 {losses[0].item():.5f}; gradient cosine {cos:.4f}")
    assert cos > 0.9, cos
