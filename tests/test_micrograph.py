"""Hand-computable micro-graph: 5 nodes, integer-valued features and weights, closed-form
radius graph / EdgeConv(max, add) / per-event MET / loss / dL/dw written out BY HAND below -- not computed by any of the
code under test.  The oracle is held to them on the CPU, the CUDA path (through the C ABI) on the GPU.

Event 0: n0 = (0, 0), n1 = (0.3, 0), n2 = (0, 0.3), n3 = (2, 2) in (eta, phi); event 1: n4 = (0, 0).  r = 0.4:
  d2(n0,n1) = d2(n0,n2) = 0.09 < 0.16 (edges), d2(n1,n2) = 0.18 (no edge), n3 and n4 see only themselves.
  edge_index (source j, target i) sorted by (i, j), self loops kept (train.py:48 loop=True):
      j: 0 1 2 0 1 0 2 3 4
      i: 0 0 0 1 1 2 2 3 4
EdgeConv(Linear(64, 32)) with W = [2 I | I], b = 1 (model/graph_met_network.py:34-38, PyG message
nn(cat[x_i, x_j - x_i])): message(j -> i) = 2 x_i + (x_j - x_i) + 1 = x_i + x_j + 1.
Features x_n = s_n * v with s = (1, 3, -2, 5, 7) and v[c] = c - 15 (integers of both signs and a zero):
  max:  out_0 = v + 1 + max(v, 3v, -2v)      out_1 = 3v + 1 + max(v, 3v)     out_2 = -2v + 1 + max(v, -2v)
        out_3 = 10v + 1                      out_4 = 14v + 1
  add:  out_0 = 3(v + 1) + (1 + 3 - 2)v = 5v + 3      out_1 = 2(3v + 1) + 4v = 10v + 2
        out_2 = 2(-2v + 1) - v = -5v + 2              out_3 = 10v + 1            out_4 = 14v + 1
loss_fn (model/net.py:49-62) with w = (1/2, 1/4, 1, 0 | 1), px = (2, 4, -2, 8 | 3), py = (0, 4, 2, 8 | -4),
truth = ((1, -1), (-3, 4)):
  MET_0 = (1 + 1 - 2 + 0, 0 + 1 + 2 + 0) = (0, 3), MET_1 = (3, -4)
  loss = 1/2 * mean((0 + 1)^2 + (3 - 1)^2, 0) = 1/2 * 5/2 = 1.25
  dL/dw_n = ((METx + tx) px_n + (METy + ty) py_n) / B = (1 * px_n + 2 * py_n) / 2 in event 0, 0 in event 1
          = (1, 6, 1, 12 | 0)
"""
import numpy as np
import pytest
import torch

from oracle import graphmet_oracle as O

ETAPHI = torch.tensor([[0.0, 0.0], [0.3, 0.0], [0.0, 0.3], [2.0, 2.0], [0.0, 0.0]])
BATCH = torch.tensor([0, 0, 0, 0, 1])
EDGE_INDEX = torch.tensor([[0, 1, 2, 0, 1, 0, 2, 3, 4], [0, 0, 0, 1, 1, 2, 2, 3, 4]])
V = np.arange(32) - 15
S = np.array([1, 3, -2, 5, 7])


def _features():
    return torch.from_numpy((S[:, None] * V[None, :]).astype(np.float32))


def _edge_mlp():
    lin = torch.nn.Linear(64, 32)
    with torch.no_grad():
        lin.weight.zero_()
        lin.weight[:, :32] = 2 * torch.eye(32)
        lin.weight[:, 32:] = torch.eye(32)
        lin.bias.fill_(1.0)
    return torch.nn.Sequential(lin)


def _expected(aggr):
    v = V.astype(np.int64)
    if aggr == "max":
        rows = [v + 1 + np.maximum.reduce([v, 3 * v, -2 * v]), 3 * v + 1 + np.maximum(v, 3 * v),
                -2 * v + 1 + np.maximum(v, -2 * v), 10 * v + 1, 14 * v + 1]
    else:
        rows = [5 * v + 3, 10 * v + 2, -5 * v + 2, 10 * v + 1, 14 * v + 1]
    return torch.from_numpy(np.stack(rows).astype(np.float32))


W = torch.tensor([0.5, 0.25, 1.0, 0.0, 1.0])
PRED = torch.zeros(5, 11)
PRED[:, 0] = torch.tensor([2.0, 4.0, -2.0, 8.0, 3.0])
PRED[:, 1] = torch.tensor([0.0, 4.0, 2.0, 8.0, -4.0])
TRUTH = torch.zeros(2, 11)
TRUTH[0, :2] = torch.tensor([1.0, -1.0])
TRUTH[1, :2] = torch.tensor([-3.0, 4.0])
LOSS = 1.25
GRAD_W = torch.tensor([1.0, 6.0, 1.0, 12.0, 0.0])


# ------------------------------------------------------------------------------------------------ oracle (CPU)
def test_oracle_radius_graph_micro():
    assert torch.equal(O.radius_graph(ETAPHI, 0.4, BATCH, loop=True, max_num_neighbors=255), EDGE_INDEX)
    no_loop = EDGE_INDEX[:, EDGE_INDEX[0] != EDGE_INDEX[1]]
    assert torch.equal(O.radius_graph(ETAPHI, 0.4, BATCH, loop=False, max_num_neighbors=255), no_loop)
    # cap 2, loop=True: target 0 keeps its two LOWEST sources (0, 1) and loses 2 [upstream: first-max-by-index]
    capped = EDGE_INDEX[:, [0, 1, 3, 4, 5, 6, 7, 8]]
    assert torch.equal(O.radius_graph(ETAPHI, 0.4, BATCH, loop=True, max_num_neighbors=2), capped)


@pytest.mark.parametrize("aggr", ["max", "add"])
def test_oracle_edgeconv_micro(aggr):
    out = O.EdgeConv(_edge_mlp(), aggr)(_features(), EDGE_INDEX)
    assert torch.equal(out.detach(), _expected(aggr))


def test_oracle_loss_micro():
    w = W.clone().requires_grad_(True)
    loss = O.loss_fn(w, PRED, TRUTH, BATCH)
    loss.backward()
    assert loss.item() == LOSS
    assert torch.equal(w.grad, GRAD_W)


# ------------------------------------------------------------------------------------------------ CUDA path (GPU)
@pytest.mark.gpu
def test_cuda_radius_graph_micro():
    from deepmetv2_b200 import ops
    pos, b = ETAPHI.cuda(), BATCH.cuda()
    assert torch.equal(ops.radius_graph(pos, 0.4, b, loop=True, max_num_neighbors=255).cpu(), EDGE_INDEX)
    no_loop = EDGE_INDEX[:, EDGE_INDEX[0] != EDGE_INDEX[1]]
    assert torch.equal(ops.radius_graph(pos, 0.4, b, loop=False, max_num_neighbors=255).cpu(), no_loop)
    capped = EDGE_INDEX[:, [0, 1, 3, 4, 5, 6, 7, 8]]
    assert torch.equal(ops.radius_graph(pos, 0.4, b, loop=True, max_num_neighbors=2).cpu(), capped)


@pytest.mark.gpu
@pytest.mark.parametrize("aggr", ["max", "add"])
@pytest.mark.parametrize("path", ["csr", "implicit", "implicit_bf16", "implicit_tf32"])
def test_cuda_edgeconv_micro(aggr, path):
    """Small integers survive bf16 / tf32 rounding, so even the tcgen05 variants must hit the closed form exactly."""
    from deepmetv2_b200 import ops
    from deepmetv2_b200.model.graph_met_network import EdgeConv
    mlp = _edge_mlp().cuda()
    x = _features().cuda()
    if path == "csr":
        ei = ops.radius_graph(ETAPHI.cuda(), 0.4, BATCH.cuda(), loop=True, max_num_neighbors=255)
        conv = EdgeConv(mlp, aggr)
        with torch.no_grad():
            out = conv(x, ei)
    else:
        ptr = torch.tensor([0, 4, 5], dtype=torch.int32, device="cuda")
        geo = ops.build_geo(ETAPHI.cuda(), 0.4, ptr, loop=True, max_num_neighbors=255)
        perm = geo.perm[:5].long()
        prec = path.split("_")[1] if "_" in path else "fp32"
        got = ops.edgeconv_geo(x[perm], geo, mlp[0].weight.detach(), mlp[0].bias.detach(), aggr=aggr, precision=prec)
        out = torch.empty_like(got)
        out[perm] = got
    assert torch.equal(out.cpu(), _expected(aggr))


@pytest.mark.gpu
def test_cuda_loss_micro():
    from deepmetv2_b200.model import net as gnet
    w = W.cuda().requires_grad_(True)
    loss = gnet.loss_fn(w, PRED.cuda(), TRUTH.cuda(), BATCH.cuda())
    loss.backward()
    assert loss.item() == LOSS
    assert torch.equal(w.grad.cpu(), GRAD_W)
