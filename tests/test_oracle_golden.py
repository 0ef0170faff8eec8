"""The oracle restatement must reproduce the fixtures produced by the reference's own
model/graph_met_network.py + model/net.py (oracle/make_golden.py) bit-for-bit, and the
reference's shipped checkpoints must strict-load."""
import os

import numpy as np
import pytest
import torch

from conftest import golden_state_dict
from oracle import graphmet_oracle as O

REF = "/root/reference"


def _inputs(g):
    x = torch.from_numpy(g["x"]); y = torch.from_numpy(g["y"]); b = torch.from_numpy(g["batch"])
    return x, y, b


def test_oracle_matches_reference_fixture(golden):
    name, g = golden
    torch.set_num_threads(1)
    x, y, b = _inputs(g)
    x_cont, x_cat, etaphi = O.prepare_inputs(x)
    assert np.array_equal(etaphi.numpy(), g["etaphi"])
    ei = O.radius_graph(etaphi, r=0.4, batch=b, loop=True, max_num_neighbors=255)
    assert np.array_equal(ei.numpy(), g["edge_index"].astype(np.int64))
    model = O.Net(8, 3)
    model.load_state_dict(golden_state_dict(g), strict=True)
    model.eval()
    w = model(x_cont, x_cat, ei, b)
    assert np.array_equal(w.detach().numpy(), g["eval_weights"])
    assert np.float32(O.loss_fn(w, x, y, b).item()) == g["eval_loss"]
    res, qT = O.metrics["resolution"](w, x, y, b)
    assert np.array_equal(qT, g["eval_qT"])
    for k, v in res.items():
        for i, nm in enumerate(("u_perp", "u_par", "R")):
            assert np.array_equal(v[i], g[f"eval_res::{k}::{nm}"]), (k, nm)
    assert np.float32(O.u_perp_par_loss(w, x, y, b).item()) == g["eval_u_perp_par_loss"]
    if "train_loss" in g:
        model.load_state_dict(golden_state_dict(g), strict=True)
        model.train(); model.zero_grad()
        w = model(x_cont, x_cat, ei, b)
        loss = O.loss_fn(w, x, y, b)
        loss.backward()
        assert np.array_equal(w.detach().numpy(), g["train_weights"])
        assert np.float32(loss.item()) == g["train_loss"]
        for k, p in model.named_parameters():
            assert np.array_equal(p.grad.numpy(), g["grad::" + k]), k
        for k, v in model.state_dict().items():
            if "running" in k or "num_batches" in k:
                assert np.array_equal(v.numpy(), g["sd_after::" + k]), k


def test_parameter_inventory():
    m = O.Net(8, 3)
    ps = list(m.parameters())
    assert len(ps) == 23 and sum(p.numel() for p in ps) == 6641
    assert len(m.state_dict()) == 32


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree only exists in the build container")
@pytest.mark.parametrize("ck", ["ckpts_dytt/best", "ckpts_dytt/last", "ckpts_znunu/best", "ckpts_znunu/last"])
def test_shipped_checkpoints_strict_load(ck):
    c = torch.load(f"{REF}/{ck}.pth.tar", map_location="cpu", weights_only=False)
    m = O.Net(8, 3)
    m.load_state_dict(c["state_dict"], strict=True)
    opt = torch.optim.AdamW(m.parameters(), lr=0.001)
    opt.load_state_dict(c["optim_dict"])
    assert len(opt.state_dict()["state"]) == 23
