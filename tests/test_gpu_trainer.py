"""The fused sync-free trainer must reproduce the reference-shaped loop (train.py:39-55 with
torch.optim.AdamW) step for step, and the oracle on CPU."""
import numpy as np
import pytest
import torch

from deepmetv2_b200 import ops, synth
from deepmetv2_b200.model import net as gnet
from deepmetv2_b200.trainer import GraphMETTrainer
from oracle import graphmet_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _dropin_step(model, opt, g):
    """train.py:40-55 verbatim, with our drop-in imports."""
    opt.zero_grad()
    x_cont = g.x[:, :8]
    x_cat = g.x[:, 8:].long()
    phi = torch.atan2(g.x[:, 1], g.x[:, 0])
    etaphi = torch.cat([g.x[:, 3][:, None], phi[:, None]], dim=1)
    edge_index = ops.radius_graph(etaphi, r=0.4, batch=g.batch, loop=True, max_num_neighbors=255)
    result = model(x_cont, x_cat, edge_index, g.batch)
    loss = gnet.loss_fn(result, g.x, g.y, g.batch)
    loss.backward()
    opt.step()
    return loss.item()


@pytest.mark.parametrize("graph", ["implicit", "csr"])
def test_trainer_matches_dropin_loop_and_oracle(graph):
    torch.manual_seed(0)
    ref = O.Net(8, 3)
    a = gnet.Net(8, 3); a.load_state_dict(ref.state_dict()); a.to(DEV)
    b = gnet.Net(8, 3); b.load_state_dict(ref.state_dict()); b.to(DEV)
    opt_ref = torch.optim.AdamW(ref.parameters(), lr=1e-3)
    opt_a = torch.optim.AdamW(a.parameters(), lr=1e-3)
    trainer = GraphMETTrainer(b, lr=1e-3, graph=graph)
    ref.train(); a.train()
    for step in range(3):
        batch = synth.make_batch(3, "dytt", seed=50 + step, sizes=[220, 260, 180])
        g = batch.to(DEV)
        la = _dropin_step(a, opt_a, g)
        lb = float(trainer.train_step(g).item())
        # oracle on CPU
        opt_ref.zero_grad()
        w, lr_, _ = O.forward_step(ref, batch.x, batch.y, batch.batch)
        lr_.backward(); opt_ref.step()
        assert abs(la - lb) <= 1e-5 * abs(la), (la, lb)
        assert abs(lb - lr_.item()) <= 1e-3 * abs(lr_.item())
    for (k, pa), (_, pb), (_, pr) in zip(a.named_parameters(), b.named_parameters(), ref.named_parameters()):
        if k.endswith("nn.0.bias"):
            # an EdgeConv bias feeds a train-mode BatchNorm: its gradient is pure rounding noise and AdamW
            # turns noise of any size into +-lr steps: each implementation random-walks by <= steps*lr, so two
            # correct implementations differ by <= 2*steps*lr (3 steps, lr 1e-3)
            assert torch.allclose(pa, pb, rtol=0, atol=6.2e-3) and torch.allclose(pb.cpu(), pr, rtol=0, atol=6.2e-3), k
            continue
        assert torch.allclose(pa, pb, rtol=1e-4, atol=1e-6), k
        assert torch.allclose(pb.cpu(), pr, rtol=2e-3, atol=2e-5), k
    for (k, va), (_, vb) in zip(a.state_dict().items(), b.state_dict().items()):
        if "running" in k or "num_batches" in k:
            # the running mean of an EdgeConv BatchNorm tracks the (random-walking, see above) bias
            atol = 6.2e-3 if "conv_continuous" in k and "running_mean" in k else 1e-6
            assert torch.allclose(va.float(), vb.float(), rtol=1e-4, atol=atol), k
    # optimizer state in torch.optim.AdamW layout
    sd = trainer.optimizer_state_dict()
    ref_sd = opt_a.state_dict()
    assert len(sd["state"]) == 23 and sd["param_groups"][0]["params"] == list(range(23))
    names = [k for k, _ in a.named_parameters()]
    for i in range(23):
        if names[i].endswith("nn.0.bias"):
            continue          # first moment of a pure-noise gradient (see above)
        ea, eb = sd["state"][i]["exp_avg"], ref_sd["state"][i]["exp_avg"]
        assert torch.allclose(ea, eb, rtol=1e-3, atol=1e-5 * float(eb.abs().max())), names[i]
    # fused eval step == drop-in eval forward
    g = synth.make_batch(2, "dytt", seed=77, sizes=[150, 210]).to(DEV)
    w, loss, met = trainer.eval_step(g)
    b.eval()
    with torch.no_grad():
        x_cont, x_cat, etaphi = ops.prepare_inputs(g.x)
        ei = ops.radius_graph(etaphi, 0.4, g.batch, loop=True, max_num_neighbors=255)
        w2 = b(x_cont, x_cat, ei, g.batch)
    assert torch.equal(w, w2)
    assert abs(loss.item() - gnet.loss_fn(w2, g.x, g.y, g.batch).item()) < 1e-6 * abs(loss.item()) + 1e-6


@pytest.mark.parametrize("graph", ["implicit", "csr"])
def test_trainer_from_host_and_checkpoint_roundtrip(tmp_path, graph):
    torch.manual_seed(1)
    m = gnet.Net(8, 3).to(DEV)
    tr = GraphMETTrainer(m, graph=graph)
    hb = synth.make_batch(2, "dytt", seed=9, sizes=[120, 140]).pin_memory()
    l0 = tr.train_step_from_host(hb)
    l1 = tr.train_step_from_host(hb)
    assert np.isfinite(l0) and np.isfinite(l1) and l1 != l0
    # parameters are still ordinary state_dict entries (utils.save_checkpoint of the reference just works)
    path = tmp_path / "last.pth.tar"
    torch.save({"epoch": 1, "state_dict": m.state_dict(), "optim_dict": tr.optimizer_state_dict()}, path)
    ck = torch.load(path, weights_only=False)
    m2 = gnet.Net(8, 3).to(DEV)
    m2.load_state_dict(ck["state_dict"])
    tr2 = GraphMETTrainer(m2, graph=graph)
    tr2.load_optimizer_state_dict(ck["optim_dict"])
    assert tr2.step_count == 2
    assert abs(tr.train_step_from_host(hb) - tr2.train_step_from_host(hb)) < 1e-6 * abs(l0)


@pytest.mark.parametrize("graph", ["implicit", "csr"])
def test_trainer_capped_graph(graph):
    """Degree cap binds -> asymmetric graph.  csr: the trainer decides on the device to use the transposed CSR;
    implicit: the cap is the per-target source-index threshold and the backward needs no transposed graph."""
    torch.manual_seed(2)
    ref = O.Net(8, 3)
    m = gnet.Net(8, 3); m.load_state_dict(ref.state_dict()); m.to(DEV)
    tr = GraphMETTrainer(m, max_num_neighbors=16, graph=graph)
    b = synth.make_batch(2, "dytt", seed=31, sizes=[260, 240])
    b.x[:, 3] *= 0.05
    b.x[:, 0] = b.x[:, 2] * torch.cos(torch.atan2(b.x[:, 1], b.x[:, 0]))   # keep (px,py) consistent
    ref.train()
    xc, xk, etaphi = O.prepare_inputs(b.x)
    # feed the oracle the kernel's own phi so the edge sets are comparable
    g = b.to(DEV)
    _, _, etaphi_dev = ops.prepare_inputs(g.x)
    ei = O.radius_graph(etaphi_dev.cpu(), 0.4, b.batch, loop=True, max_num_neighbors=16)
    w = ref(xc, xk, ei, b.batch)
    loss = O.loss_fn(w, b.x, b.y, b.batch)
    loss.backward()
    lt = tr.train_step(g)
    assert tr.graph_stats()[2] > 0                     # some nodes were capped
    assert abs(lt.item() - loss.item()) <= 1e-3 * abs(loss.item())
    gref = torch.cat([p.grad.flatten() for p in ref.parameters()])
    err = (tr.flat_grad.cpu() - gref).abs().max() / gref.abs().max()
    assert err < 1e-3, err


def test_lr_scheduler_drives_the_fused_step():
    """train.py:58,76: ReduceLROnPlateau on the trainer's optimizer handle changes what the fused AdamW kernel applies."""
    torch.manual_seed(0)
    b = synth.make_batch(2, "dytt", seed=5, sizes=[200, 180]).to(DEV)
    def run(schedule):
        m = gnet.Net(8, 3)
        torch.manual_seed(1)
        for p in m.parameters():
            torch.nn.init.normal_(p, std=0.1) if p.dim() > 1 else None
        m.to(DEV)
        tr = GraphMETTrainer(m, lr=1e-3)
        sched = torch.optim.lr_scheduler.ReduceLROnPlateau(tr.optimizer, factor=0.5, patience=0, threshold=0.05)
        before = tr.flat.clone()
        if schedule:
            sched.step(1.0); sched.step(1.0)          # no improvement with patience 0: lr halves
            assert abs(tr.lr - 5e-4) < 1e-12
        tr.train_step(b)
        torch.cuda.synchronize()
        return (tr.flat - before).abs().max().item(), tr
    d_full, _ = run(False)
    d_half, tr = run(True)
    assert abs(d_half / d_full - 0.5) < 0.02            # the first AdamW step moves every weight by ~lr
    sd = tr.optimizer.state_dict()
    assert sd["param_groups"][0]["lr"] == tr.lr and len(sd["state"]) == 23
    ref_opt = torch.optim.AdamW([torch.nn.Parameter(torch.zeros_like(v)) for v in tr._views], lr=1e-3)
    ref_opt.load_state_dict(sd)                          # torch's own AdamW accepts the layout


def test_optimiser_in_the_backward_equals_the_two_call_form(monkeypatch):
    """gm_net_backward_geo_adamw (AdamW applied by the last CTA of the backward's final reduction) must leave the
    parameters, the optimiser state and the status mirror BIT-identical to gm_net_backward_geo + gm_adamw_step_ex."""
    from deepmetv2_b200 import trainer as T
    batches = [synth.make_batch(6, "dytt", seed=70 + i).to(DEV) for i in range(3)]
    states = []
    for fused in (True, False):
        monkeypatch.setattr(T, "_FUSED_ADAMW", fused)
        torch.manual_seed(5)
        tr = GraphMETTrainer(gnet.Net(8, 3).to(DEV), lr=2e-3, weight_decay=0.01)
        losses = [tr.train_step(b).clone() for b in batches for _ in range(2)]
        torch.cuda.synchronize()
        states.append((tr.flat.clone(), tr.exp_avg.clone(), tr.exp_avg_sq.clone(), torch.stack(losses), tr.step_count))
        tr.close()
    for a, b in zip(states[0][:4], states[1][:4]):
        assert torch.equal(a, b)
    assert states[0][4] == states[1][4] == 6
