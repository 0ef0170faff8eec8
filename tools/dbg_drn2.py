import sys, types, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import numpy as np, torch
from test_gpu_drn import *
z, sd = load_case("small")
ref = D.DynamicReductionNetwork(input_dim=11, hidden_dim=64, k=8, output_dim=2, aggr="max", norm=torch.tensor(D.DRN_NORM)); ref.load_state_dict(sd)
ours = drn.DynamicReductionNetwork(input_dim=11, hidden_dim=64, k=8, output_dim=2, aggr="max", norm=torch.tensor(D.DRN_NORM)); ours.load_state_dict(sd); ours.to(DEV)
ref.train(); ours.train()
x, batch = torch.from_numpy(z["x"]), torch.from_numpy(z["batch"])
B = int(batch.max()) + 1; N = x.shape[0]
R = torch.randn(B, 2, generator=torch.Generator().manual_seed(7))
out = ours(types.SimpleNamespace(x=x.to(DEV), batch=batch.to(DEV)))
(out * R.to(DEV)).sum().backward(); torch.cuda.synchronize()
it = {k: v.cpu() for k, v in ours.intermediates().items()}
ei1, ei2 = csr_to_ei(it["rowptr1"], it["col1"]), csr_to_ei(it["rowptr2"], it["col2"])
# oracle, stage by stage, with leaves we can read gradients from
x0 = ref.inputnet(ref.datanorm * x); x0.retain_grad()
arg1, arg2 = (t.cpu() for t in ours.selected_sources())
from test_gpu_drn import _edgeconv_selected
x1 = _edgeconv_selected(ref.edgeconv1, x0, ei1, arg1); x1.retain_grad()
cl = it["cluster1"].long(); newid, _ = D.consecutive_cluster(cl); M = int(newid.max()) + 1
lo = torch.full((M,), N, dtype=torch.long).scatter_reduce(0, newid, torch.arange(N), "amin")
hi = torch.zeros(M, dtype=torch.long).scatter_reduce(0, newid, torch.arange(N), "amax")
win = torch.where(it["x1"][lo] >= it["x1"][hi], lo[:, None].expand(-1, 64), hi[:, None].expand(-1, 64))
xp = x1.gather(0, win); bp = batch[lo]; xp.retain_grad()
x2 = _edgeconv_selected(ref.edgeconv2, xp, ei2, arg2[:xp.shape[0]]); x2.retain_grad()
o = ref.output(D.global_max_pool(x2, bp)).squeeze(-1)
(o * R).sum().backward()
# our backward workspace (mirrors drn_bwd_layout)
al = lambda v: (v + 255) // 256 * 256
K = 9; ecap = 2 * N * K
off = 0
def take(nbytes):
    global off
    o_ = off; off = al(off + nbytes); return o_
o_dz1 = take(ecap * 96 * 4); o_mir = take(ecap * 4); o_dab = take(N * 192 * 4); o_gA = take(N * 64 * 4); o_gB = take(N * 64 * 4)
bws = ours._bws
f32 = lambda o_, n: bws[o_:o_ + n * 4].view(torch.float32).cpu()
gA = f32(o_gA, N * 64).view(N, 64); gB = f32(o_gB, N * 64).view(N, 64)
rel = lambda a, b: float((a - b).abs().max() / b.abs().max())
print("dx1 (gA) vs oracle:", rel(gA, x1.grad))
print("dx0 (gB) vs oracle:", rel(gB, x0.grad))
d = (gB - x0.grad).abs().max(1).values
print("rows with dx0 error > 1e-4*max:", int((d > 1e-4 * x0.grad.abs().max()).sum()), "of", N, "worst rows", torch.topk(d, 8).indices.tolist())
d1 = (gA - x1.grad).abs()
print("dx1 mismatching entries:", int((d1 > 1e-5 * x1.grad.abs().max()).sum()))
bad = (d1 > 1e-5 * x1.grad.abs().max()).nonzero()[:10]
for r, c in bad.tolist():
    print("  node", r, "ch", c, "ours", float(gA[r, c]), "oracle", float(x1.grad[r, c]), "cluster", int(it["cluster1"][r]), "partner", int(it["partner1"][r]))

dab = f32(o_dab, N * 192).view(N, 192)
# oracle dA, dB: A = x0 (W1a - W1b)^T + b1, B = x0 W1b^T: recompute with leaves
lin1 = ref.edgeconv1.nn[0]
W1a, W1b = lin1.weight[:, :64], lin1.weight[:, 64:]
x0d = x0.detach()
A = (x0d @ (W1a - W1b).t() + lin1.bias).detach().requires_grad_(); Bv = (x0d @ W1b.t()).detach().requires_grad_()
j, i = ei1[0], ei1[1]
import torch.nn.functional as F
h1 = F.elu(A[i] + Bv[j]); h2 = F.elu(ref.edgeconv1.nn[2](h1)); h2.retain_grad()
ref.edgeconv1.nn[4].train()
msg = ref.edgeconv1.nn[4](h2)
x1b = msg.new_zeros(N, 64).scatter_reduce(0, i.view(-1, 1).expand(-1, 64), msg, "amax", include_self=False)
x1b.backward(x1.grad)
print("dA vs oracle:", rel(dab[:, :96], A.grad), " dB vs oracle:", rel(dab[:, 96:], Bv.grad))
dA_err = (dab[:, :96] - A.grad).abs().max(1).values; dB_err = (dab[:, 96:] - Bv.grad).abs().max(1).values
print("rows dA bad:", int((dA_err > 1e-4 * A.grad.abs().max()).sum()), torch.topk(dA_err, 5).indices.tolist(), " rows dB bad:", int((dB_err > 1e-4 * Bv.grad.abs().max()).sum()), torch.topk(dB_err, 5).indices.tolist())
# per-edge dz2 check through ties: which (target, channel) have several edges at the max
cnt = torch.zeros(N, 64)
sel = (msg == x1b[i]).float()
cnt.index_add_(0, i, sel)
print("oracle-side (target, channel) with tied maxima:", int((cnt > 1).sum()), "max multiplicity", int(cnt.max()))
rows = (cnt > 1).nonzero()[:8].tolist(); print(rows)
E1 = int(it["rowptr1"][-1])
mir = bws[o_mir:o_mir + E1 * 4].view(torch.int32).cpu().long()
row1, col1 = it["row1"].long(), it["col1"].long()
okm = (col1[mir] == row1) & (row1[mir] == col1)
print("mirror ok:", bool(okm.all()), "bad entries:", int((~okm).sum()), (~okm).nonzero()[:5].flatten().tolist())
dz1 = f32(o_dz1, E1 * 96).view(E1, 96)
# oracle per-edge dz1 = d/d(A_i + B_j)
dz1_ref = torch.autograd.grad  # placeholder
A2 = A.detach().clone().requires_grad_(); B2 = Bv.detach().clone().requires_grad_()
z1 = (A2[i] + B2[j]); z1.retain_grad()
h1b = F.elu(z1); h2b = F.elu(ref.edgeconv1.nn[2](h1b)); msgb = ref.edgeconv1.nn[4](h2b)
keyb = i * N + j; orderb = torch.argsort(keyb)
qb = torch.arange(N)[:, None] * N + arg1.long().clamp_min(0)
eidxb = orderb[torch.searchsorted(keyb[orderb], qb.reshape(-1)).clamp_max(keyb.numel() - 1)].view(N, 64)
x1c = msgb.gather(0, eidxb)
x1c.backward(x1.grad)
# edge order of ei1: sorted by (row = ei1[0]); our CSR entry p = (row r, col c) is the edge with TARGET r and SOURCE c
# ei1[0] = row (we treat as source j above!), ei1[1] = col (target i)
print("edge p: ours target row1[p], source col1[p]; oracle used j = ei1[0], i = ei1[1]")
key_ours = row1 * N + col1      # (target, source)
key_ref = i * N + j             # (target, source) of oracle edge list
perm = torch.argsort(key_ref); assert torch.equal(key_ref[perm], torch.sort(key_ours).values)
ours_sorted = dz1[torch.argsort(key_ours)]; ref_sorted = z1.grad[perm]
e_err = (ours_sorted - ref_sorted).abs().max(1).values
print("dz1 vs oracle:", float(e_err.max() / ref_sorted.abs().max()), "bad edges:", int((e_err > 1e-4 * ref_sorted.abs().max()).sum()))

ko = torch.argsort(key_ours); worst = torch.topk(e_err, 6).indices
for w_ in worst.tolist():
    p = int(ko[w_]); print("edge", p, "target", int(row1[p]), "source", int(col1[p]), "err", float(e_err[w_]), "tile", p // 128, "pos", p % 128)
