import sys, types, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import numpy as np, torch
from test_gpu_drn import *
z, sd = load_case("small")
ref = D.DynamicReductionNetwork(input_dim=11, hidden_dim=64, k=8, output_dim=2, aggr="max", norm=torch.tensor(D.DRN_NORM)); ref.load_state_dict(sd)
ours = drn.DynamicReductionNetwork(input_dim=11, hidden_dim=64, k=8, output_dim=2, aggr="max", norm=torch.tensor(D.DRN_NORM)); ours.load_state_dict(sd); ours.to(DEV)
ref.train(); ours.train()
x, batch = torch.from_numpy(z["x"]), torch.from_numpy(z["batch"])
B = int(batch.max()) + 1
R = torch.randn(B, 2, generator=torch.Generator().manual_seed(7))
out = ours(types.SimpleNamespace(x=x.to(DEV), batch=batch.to(DEV)))
(out * R.to(DEV)).sum().backward(); torch.cuda.synchronize()
it = {k: v.cpu() for k, v in ours.intermediates().items()}
graphs = {1: csr_to_ei(it["rowptr1"], it["col1"]), 2: csr_to_ei(it["rowptr2"], it["col2"])}
hooks = {"graph": lambda s, xx, bb: graphs[s], "cluster": lambda s, ei, w, n: it["cluster1"].long() if s == 1 else D.graclus(ei, w.detach(), n)}
out_ref = ref(types.SimpleNamespace(x=x, batch=batch), hooks)
(out_ref * R).sum().backward()
named = dict(ours.named_parameters())
for k, p in ref.named_parameters():
    g, gr = named[k].grad.cpu(), p.grad
    print(f"{k:28s} err {float((g-gr).abs().max()/gr.abs().max().clamp_min(1e-30)):.3e}  max {float(gr.abs().max()):.3e}")
print("db2 stage1 sum (matches if debug build):", float(named["edgeconv1.nn.2.bias"].grad.sum()), "stage2:", float(named["edgeconv2.nn.2.bias"].grad.sum()))
deg1 = (it["rowptr1"][1:] - it["rowptr1"][:-1]); deg2 = (it["rowptr2"][1:] - it["rowptr2"][:-1])
print("expected", int((deg1 > 0).sum()) * 64, int((deg2 > 0).sum()) * 64)
print("per-channel stage1:", named["edgeconv1.nn.2.bias"].grad.cpu().int().tolist())
g1 = {k: v.grad.clone() for k, v in named.items()}
ours.zero_grad()
out2 = ours(types.SimpleNamespace(x=x.to(DEV), batch=batch.to(DEV)))
(out2 * R.to(DEV)).sum().backward(); torch.cuda.synchronize()
print("run-to-run identical:", all(torch.equal(named[k].grad, g1[k]) for k in g1))
