"""Print per-tensor gradient errors of the CUDA path vs the committed golden fixtures (GPU box)."""
import sys, os
import numpy as np, torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO); sys.path.insert(0, os.path.join(REPO, "tests"))
from conftest import load_golden, golden_state_dict
from deepmetv2_b200 import ops
from deepmetv2_b200.model import net as gnet

def nerr(a, b, floor=1e-12):
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), floor))

for name in ["tiny_init", "tiny_dytt_best", "small_init", "small_znunu_last"]:
    g = load_golden(name)
    m = gnet.Net(8, 3); m.load_state_dict(golden_state_dict(g)); m.to("cuda"); m.train()
    if len(sys.argv) > 1:
        m.graphnet.edge_precision = sys.argv[1]
    x = torch.from_numpy(g["x"]).cuda(); y = torch.from_numpy(g["y"]).cuda(); b = torch.from_numpy(g["batch"]).cuda()
    ei = ops.radius_graph(torch.from_numpy(g["etaphi"]).cuda(), r=0.4, batch=b, loop=True, max_num_neighbors=255)
    w = m(x[:, :8], x[:, 8:].long(), ei, b)
    loss = gnet.loss_fn(w, x, y, b); loss.backward()
    print(name, "N", x.shape[0], "E", ei.shape[1], "w rel", float(np.max(np.abs(w.detach().cpu().numpy() - g["train_weights"]) / np.abs(g["train_weights"]))))
    for k, p in m.named_parameters():
        print(f"   {k:50s} {nerr(p.grad.cpu().numpy(), g['grad::' + k]):.3e}  max|g| {np.abs(g['grad::' + k]).max():.3e}")
