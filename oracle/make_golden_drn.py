"""ORACLE — TEST INFRASTRUCTURE ONLY.  Generates tests/golden/golden_drn_*.npz.

Runs in the build container only.  Imports the reference's OWN /root/reference/model/dynamic_reduction_network.py
unmodified (the absent upstream packages are replaced by oracle/upstream_stubs.py, whose DRN operators are the
restatements in oracle/drn_oracle.py), builds it with the arguments of the commented wrapper at model/net.py:14-29
(input_dim 11, hidden 64, k 8, output_dim 2, aggr 'max'), runs ``forward(data)`` in eval and train mode on small seeded
batches and records inputs, parameters, every intermediate the stubs see and the outputs.

    python -m oracle.make_golden_drn            # from the repo root
"""
import os
import sys
import types

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
sys.path.insert(0, REPO)

from oracle import upstream_stubs  # noqa: E402
from oracle import drn_oracle as D  # noqa: E402
from deepmetv2_b200 import synth  # noqa: E402


def _import_reference():
    upstream_stubs.install()
    sys.path.insert(0, REF)
    import model.dynamic_reduction_network as ref_drn
    assert ref_drn.__file__.startswith(REF), ref_drn.__file__
    return ref_drn


def _recording(ref_drn, rec):
    """Wrap the stub operators the reference file bound at import so each call's result lands in ``rec``."""
    saved = {n: getattr(ref_drn, n) for n in ("knn_graph", "to_undirected", "normalized_cut", "graclus", "max_pool")}
    count = {"g": 0, "w": 0, "c": 0}

    def to_undirected(ei, *a, **k):
        out = saved["to_undirected"](ei, *a, **k)
        count["g"] += 1
        rec[f"edge_index{count['g']}"] = out.numpy().astype(np.int32)
        return out

    def normalized_cut(ei, attr, num_nodes=None):
        out = saved["normalized_cut"](ei, attr, num_nodes)
        count["w"] += 1
        rec[f"weight{count['w']}"] = out.detach().numpy()
        return out

    def graclus(ei, w, n):
        out = saved["graclus"](ei, w, n)
        count["c"] += 1
        rec[f"cluster{count['c']}"] = out.numpy().astype(np.int32)
        return out

    def max_pool(cluster, data):
        out = saved["max_pool"](cluster, data)
        rec["x1p"] = out.x.detach().numpy()
        rec["batch1p"] = out.batch.numpy().astype(np.int32)
        return out

    ref_drn.to_undirected, ref_drn.normalized_cut = to_undirected, normalized_cut
    ref_drn.graclus, ref_drn.max_pool = graclus, max_pool
    return saved


def run_case(ref_drn, name, batch, out_dir, seed):
    torch.manual_seed(seed)
    model = ref_drn.DynamicReductionNetwork(input_dim=11, hidden_dim=64, k=8, output_dim=2, aggr="max",
                                            norm=torch.tensor(D.DRN_NORM))
    # non-trivial BatchNorm state (default init would make eval-mode BN the identity)
    with torch.no_grad():
        for ec in (model.edgeconv1, model.edgeconv2):
            bn = ec.nn[4]
            bn.weight.uniform_(-1.5, 1.5)          # negative gammas: max_j BN(h_j) is then BN(min_j h_j)
            bn.bias.uniform_(-0.5, 0.5)
            bn.running_mean.uniform_(-0.2, 0.4)
            bn.running_var.uniform_(0.5, 1.5)
    sd0 = {k: v.clone() for k, v in model.state_dict().items()}
    rec = {"x": batch.x.numpy(), "batch": batch.batch.numpy(), "ptr": batch.ptr.numpy()}
    rec.update({"sd::" + k: v.numpy() for k, v in sd0.items()})
    for mode in ("eval", "train"):
        model.load_state_dict(sd0)
        model.train(mode == "train")
        sub = {}
        saved = _recording(ref_drn, sub)
        hooks = [model.inputnet.register_forward_hook(lambda m, i, o: sub.__setitem__("x0", o.detach().numpy())),
                 model.edgeconv1.register_forward_hook(lambda m, i, o: sub.__setitem__("x1", o.detach().numpy())),
                 model.edgeconv2.register_forward_hook(lambda m, i, o: sub.__setitem__("x2", o.detach().numpy()))]
        data = types.SimpleNamespace(x=batch.x.clone(), batch=batch.batch.clone(), edge_index=None, edge_attr=None,
                                     pos=None)
        out = model(data)                      # dynamic_reduction_network.py:81-104, verbatim
        for h in hooks:
            h.remove()
        for n, f in saved.items():
            setattr(ref_drn, n, f)
        sub["out"] = out.detach().numpy()
        # the commented wrapper, model/net.py:31-36
        met = torch.nn.Softplus()(out[:, 0]).unsqueeze(1)
        metphi = np.pi * (2 * torch.sigmoid(out[:, 1]) - 1).unsqueeze(1)
        sub["net_out"] = torch.cat((met, metphi), 1).detach().numpy()
        if mode == "train":
            sub.update({"sd_after::" + k: v.numpy() for k, v in model.state_dict().items()
                        if "running" in k or "num_batches" in k})
        rec.update({f"{mode}::{k}": v for k, v in sub.items()})
    path = os.path.join(out_dir, f"golden_drn_{name}.npz")
    np.savez_compressed(path, **rec)
    print(f"drn_{name}: N={batch.x.shape[0]} B={batch.y.shape[0]} E1={rec['eval::edge_index1'].shape[1]} "
          f"M={rec['eval::x1p'].shape[0]} E2={rec['eval::edge_index2'].shape[1]} -> {os.path.relpath(path, REPO)} "
          f"({os.path.getsize(path) / 1024:.0f} KiB)")


def main():
    ref_drn = _import_reference()
    out_dir = os.path.join(REPO, "tests", "golden")
    torch.set_num_threads(1)
    run_case(ref_drn, "tiny", synth.make_batch(4, "dytt", seed=21, sizes=[60, 1, 45, 30]), out_dir, 4321)
    run_case(ref_drn, "small", synth.make_batch(3, "dytt", seed=22, sizes=[300, 420, 260]), out_dir, 4322)


if __name__ == "__main__":
    main()
