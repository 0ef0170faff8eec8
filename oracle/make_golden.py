"""ORACLE — TEST INFRASTRUCTURE ONLY.  Generates tests/golden/*.npz.

Runs in the build container only (needs /root/reference; the GPU box never runs this).
It imports the reference's OWN model code — /root/reference/model/graph_met_network.py and
/root/reference/model/net.py, unmodified — with only the three absent upstream packages
replaced by oracle/upstream_stubs.py, executes train.py:42-51 / evaluate.py:83-103 on small
seeded synthetic batches and records inputs, weights and every output.  The restatement in
oracle/graphmet_oracle.py and the CUDA path are then both checked against these files.

    python -m oracle.make_golden            # from the repo root
"""
import os
import sys

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
sys.path.insert(0, REPO)

from oracle import upstream_stubs  # noqa: E402
from deepmetv2_b200 import synth  # noqa: E402


def _import_reference():
    upstream_stubs.install()
    sys.path.insert(0, REF)
    import model.net as ref_net  # the reference's file, imported in place
    assert ref_net.__file__.startswith(REF), ref_net.__file__
    from torch_cluster import radius_graph  # the stub
    return ref_net, radius_graph


def _state_to_np(sd):
    return {"sd::" + k: v.detach().cpu().numpy() for k, v in sd.items()}


def run_case(ref_net, radius_graph, name, batch, weights, out_dir, with_grads=True):
    torch.manual_seed(1234)
    model = ref_net.Net(8, 3)
    if weights != "init":
        ck = torch.load(os.path.join(REF, weights), map_location="cpu", weights_only=False)
        model.load_state_dict(ck["state_dict"])
    sd0 = {k: v.clone() for k, v in model.state_dict().items()}
    x, y, b = batch.x, batch.y, batch.batch
    rec = {"x": x.numpy(), "y": y.numpy(), "batch": b.numpy(), "ptr": batch.ptr.numpy()}
    rec.update(_state_to_np(sd0))

    # --- train.py:42-48 / evaluate.py:83-88 (verbatim op sequence) ---
    x_cont = x[:, :8]
    x_cat = x[:, 8:].long()
    phi = torch.atan2(x[:, 1], x[:, 0])
    etaphi = torch.cat([x[:, 3][:, None], phi[:, None]], dim=1)
    edge_index = radius_graph(etaphi, r=0.4, batch=b, loop=True, max_num_neighbors=255)
    rec["etaphi"] = etaphi.numpy()
    rec["edge_index"] = edge_index.numpy().astype(np.int32)

    # --- evaluate.py:42,89,100,103 ---
    model.eval()
    w = model(x_cont, x_cat, edge_index, b)
    loss = ref_net.loss_fn(w, x, y, b)
    res, qT = ref_net.metrics["resolution"](w, x, y, b)
    rec["eval_weights"] = w.detach().numpy()
    rec["eval_loss"] = np.float32(loss.item())
    rec["eval_qT"] = qT
    for k, v in res.items():
        for i, nm in enumerate(("u_perp", "u_par", "R")):
            rec[f"eval_res::{k}::{nm}"] = v[i]
    rec["eval_u_perp_par_loss"] = np.float32(ref_net.u_perp_par_loss(w, x, y, b).item())

    # --- train.py:35,49-51 ---
    if with_grads:
        model.load_state_dict(sd0)
        model.train()
        model.zero_grad()
        w = model(x_cont, x_cat, edge_index, b)
        loss = ref_net.loss_fn(w, x, y, b)
        loss.backward()
        rec["train_weights"] = w.detach().numpy()
        rec["train_loss"] = np.float32(loss.item())
        for k, p in model.named_parameters():
            rec["grad::" + k] = p.grad.detach().numpy()
        rec.update({"sd_after::" + k: v.detach().numpy() for k, v in model.state_dict().items()
                    if "running" in k or "num_batches" in k})
    path = os.path.join(out_dir, f"golden_{name}.npz")
    np.savez_compressed(path, **rec)
    print(f"{name}: N={x.shape[0]} B={y.shape[0]} E={edge_index.shape[1]} "
          f"eval_loss={rec['eval_loss']:.4f} -> {os.path.relpath(path, REPO)} "
          f"({os.path.getsize(path) / 1024:.0f} KiB)")


def main():
    ref_net, radius_graph = _import_reference()
    out_dir = os.path.join(REPO, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    torch.set_num_threads(1)  # fixed summation order in the fixtures
    tiny = synth.make_batch(5, "tiny", seed=11)
    small = synth.make_batch(3, "dytt", seed=12, sizes=[300, 420, 260])
    c1 = synth.make_batch(1, "dytt", seed=13, sizes=[1700])
    run_case(ref_net, radius_graph, "tiny_init", tiny, "init", out_dir)
    run_case(ref_net, radius_graph, "tiny_dytt_best", tiny, "ckpts_dytt/best.pth.tar", out_dir)
    run_case(ref_net, radius_graph, "small_init", small, "init", out_dir)
    run_case(ref_net, radius_graph, "small_znunu_last", small, "ckpts_znunu/last.pth.tar", out_dir)
    run_case(ref_net, radius_graph, "c1_dytt_best", c1, "ckpts_dytt/best.pth.tar", out_dir,
             with_grads=False)


if __name__ == "__main__":
    main()
