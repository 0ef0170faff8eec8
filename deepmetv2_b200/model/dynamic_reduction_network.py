"""``DynamicReductionNetwork`` with the reference's constructor / forward signature and state_dict layout
(/root/reference/model/dynamic_reduction_network.py:32-104) and the ``Net`` wrapper the reference keeps commented out at
/root/reference/model/net.py:12-37 (``DRNNet`` here), executed by libgraphmet.so (``gm_drn_forward``, csrc/drn_net.cu +
csrc/drn_graph.cu + csrc/knn.cu).

The module tree (``datanorm``, ``inputnet.{0,2,4}``, ``edgeconv{1,2}.nn.{0,2,4}``, ``output.{0,2,4}``) is the
reference's, so a checkpoint of the reference model strict-loads.  ``forward(data)`` takes any object with ``.x`` [N,11]
and ``.batch`` [N] (sorted, PyG collation) or ``.ptr`` [B+1]; it does not execute the sub-modules but hands their
storages to ONE C call that runs the whole forward without a host synchronisation; in training mode with autograd
enabled the call is one ``torch.autograd.Function`` whose backward is ONE C call (``gm_drn_backward``: all 31 gradients,
fp32).  There is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import torch
from torch import nn

from .. import _lib, ops
from .._lib import GmDrnConv, GmDrnParams, GmDrnView, check, lib, ptr, stream


class EdgeConvMLP(nn.Module):
    """Holder with torch_geometric.nn.EdgeConv's attributes (``nn``, ``aggr``, ``flow``) for the DRN edge MLP."""

    def __init__(self, nn, aggr="add"):  # noqa: A002  (upstream argument name)
        super().__init__()
        self.nn = nn
        self.aggr = aggr
        self.flow = "source_to_target"


class _DRNFunction(torch.autograd.Function):
    """One autograd node for the whole network (model/dynamic_reduction_network.py:81-104)."""

    @staticmethod
    def forward(ctx, net, x, ptr_i32, men, *params):
        out, _ = net._launch(x, ptr_i32, men, want_net_out=False)
        ctx.net, ctx.x, ctx.ptr = net, x, ptr_i32
        ctx.ws_tag = net._ws_tag
        return out

    @staticmethod
    def backward(ctx, grad_out):
        net = ctx.net
        if net._ws_tag != ctx.ws_tag:
            raise RuntimeError("DynamicReductionNetwork: another forward ran before this backward (one workspace per module)")
        params = list(net.parameters())
        dev = ctx.x.device
        flat = torch.empty(sum(q.numel() for q in params), dtype=torch.float32, device=dev)
        views, off = [], 0
        for q in params:
            views.append(flat[off:off + q.numel()].view_as(q))
            off += q.numel()
        G = net._param_struct(grad_views=dict(zip((n for n, _ in net.named_parameters()), views)))
        P = net._param_struct()
        l = lib()
        N, B = ctx.x.shape[0], ctx.ptr.numel() - 1
        need = l.gm_drn_bwd_ws_bytes(N, B, int(net.k))
        if net._bws is None or net._bws.numel() < need or net._bws.device != dev:
            net._bws = torch.empty(need, dtype=torch.uint8, device=dev)
        g = grad_out.contiguous().float().view(B, net.output_dim)
        check(l.gm_drn_backward(C.byref(P), C.byref(G), ptr(ctx.x), ctx.x.stride(0), ptr(ctx.ptr), N, B, ptr(net._ws),
                                net._ws.numel(), ptr(net._bws), net._bws.numel(), ptr(g), stream(dev)), "gm_drn_backward")
        return (None, None, None, None) + tuple(v if q.requires_grad else None for v, q in zip(views, params))


class DynamicReductionNetwork(nn.Module):
    def __init__(self, input_dim=5, hidden_dim=64, output_dim=1, k=16, aggr="add",
                 norm=torch.tensor([1. / 500., 1. / 500., 1. / 54., 1 / 25., 1. / 1000.])):
        super().__init__()
        self.datanorm = nn.Parameter(norm)
        self.k = k
        start_width = 2 * hidden_dim
        middle_width = 3 * hidden_dim // 2
        self.inputnet = nn.Sequential(
            nn.Linear(input_dim, hidden_dim // 2), nn.ELU(),
            nn.Linear(hidden_dim // 2, hidden_dim), nn.ELU(),
            nn.Linear(hidden_dim, hidden_dim), nn.ELU())

        def convnn():
            return nn.Sequential(nn.Linear(start_width, middle_width), nn.ELU(),
                                 nn.Linear(middle_width, hidden_dim), nn.ELU(),
                                 nn.BatchNorm1d(num_features=hidden_dim))
        self.edgeconv1 = EdgeConvMLP(nn=convnn(), aggr=aggr)
        self.edgeconv2 = EdgeConvMLP(nn=convnn(), aggr=aggr)
        self.output = nn.Sequential(nn.Linear(hidden_dim, hidden_dim), nn.ELU(),
                                    nn.Linear(hidden_dim, hidden_dim // 2), nn.ELU(),
                                    nn.Linear(hidden_dim // 2, output_dim))
        self.input_dim, self.hidden_dim, self.output_dim = input_dim, hidden_dim, output_dim
        self.edge_precision = "fp32"
        self._ws = None
        self._bws = None
        self._ws_tag = 0
        self._view = None
        self._sizes = (0, 0)

    # ------------------------------------------------------------------------------------------ C structs
    def _param_struct(self, grad_views: dict | None = None) -> GmDrnParams:
        """Parameter pointers, or (``grad_views``: name -> tensor) the gradient buffers in the same struct."""
        g = grad_views
        pw = (lambda mod, name, key: ptr(getattr(mod, name))) if g is None else (lambda mod, name, key: ptr(g[key]))
        if self.input_dim != 11 or self.hidden_dim != 64 or self.edgeconv1.aggr != "max" or self.output_dim > 8:
            raise NotImplementedError("the sm_100a DRN kernels are specialised for the configuration of "
                                      "model/net.py:14-29: input_dim=11, hidden_dim=64, aggr='max'")
        P = GmDrnParams()
        P.datanorm = pw(self, "datanorm", "datanorm")
        for i, idx in enumerate((0, 2, 4)):
            P.in_w[i], P.in_b[i] = pw(self.inputnet[idx], "weight", f"inputnet.{idx}.weight"), pw(self.inputnet[idx], "bias", f"inputnet.{idx}.bias")
            P.out_w[i], P.out_b[i] = pw(self.output[idx], "weight", f"output.{idx}.weight"), pw(self.output[idx], "bias", f"output.{idx}.bias")
        for i, (nm, ec) in enumerate((("edgeconv1", self.edgeconv1), ("edgeconv2", self.edgeconv2))):
            c, seq = GmDrnConv(), ec.nn
            c.w1, c.b1 = pw(seq[0], "weight", f"{nm}.nn.0.weight"), pw(seq[0], "bias", f"{nm}.nn.0.bias")
            c.w2, c.b2 = pw(seq[2], "weight", f"{nm}.nn.2.weight"), pw(seq[2], "bias", f"{nm}.nn.2.bias")
            bn = seq[4]
            c.bn_weight, c.bn_bias = pw(bn, "weight", f"{nm}.nn.4.weight"), pw(bn, "bias", f"{nm}.nn.4.bias")
            if g is None:
                c.bn_running_mean, c.bn_running_var = ptr(bn.running_mean), ptr(bn.running_var)
                c.bn_num_batches_tracked = ptr(bn.num_batches_tracked)
            P.conv[i] = c
        P.output_dim, P.k = int(self.output_dim), int(self.k)
        return P

    # ------------------------------------------------------------------------------------------ forward
    def _prep(self, data, max_event_nodes=None):
        x = data.x
        _lib.require_cuda(x, "data.x")
        if x.dtype != torch.float32 or x.dim() != 2 or x.shape[1] < 11 or x.stride(1) != 1:
            raise ValueError("data.x must be float32 [N, >=11] with unit column stride")
        p = getattr(data, "ptr", None)
        if p is None or p.dtype != torch.int32:
            batch = data.batch
            B = int(batch[-1].item()) + 1
            p = ops.batch_to_ptr(batch, B)
        if max_event_nodes is None:
            max_event_nodes = getattr(data, "max_event_nodes", None)
        if max_event_nodes is None:
            max_event_nodes = int((p[1:] - p[:-1]).max().item())
        return x, p, int(max_event_nodes)

    def _launch(self, x, p, max_event_nodes, want_net_out: bool):
        B, N = p.numel() - 1, x.shape[0]
        l, dev = lib(), x.device
        bn = self.edgeconv1.nn[4]
        need = l.gm_drn_ws_bytes(N, B, int(self.k))
        if self._ws is None or self._ws.numel() < need or self._ws.device != dev:
            self._ws = torch.empty(need, dtype=torch.uint8, device=dev)
        out = torch.empty(B, self.output_dim, dtype=torch.float32, device=dev)
        net_out = torch.empty(B, 2, dtype=torch.float32, device=dev) if want_net_out else None
        view = GmDrnView()
        P = self._param_struct()
        with torch.no_grad():
            check(l.gm_drn_forward(C.byref(P), ptr(x), x.stride(0), ptr(p), N, B, int(max_event_nodes),
                                   int(self.training), _lib.PRECISION[self.edge_precision],
                                   float(bn.momentum if bn.momentum is not None else 0.1), float(bn.eps), ptr(self._ws),
                                   self._ws.numel(), ptr(out), ptr(net_out), C.byref(view), stream(dev)), "gm_drn_forward")
        self._view, self._sizes = view, (N, B)
        self._ws_tag += 1
        return out, net_out

    def _wants_grad(self) -> bool:
        return torch.is_grad_enabled() and any(q.requires_grad for q in self.parameters())

    def forward(self, data, max_event_nodes=None):
        x, p, men = self._prep(data, max_event_nodes)
        # the backward differentiates the train-mode (batch-statistics) fp32 forward; eval mode and the bf16 tcgen05
        # variant are inference paths: their output carries no autograd graph
        if self._wants_grad() and self.training and self.edge_precision == "fp32":
            out = _DRNFunction.apply(self, x, p, men, *self.parameters())
        else:
            out, _ = self._launch(x, p, men, want_net_out=False)
        return out.squeeze(-1)

    # ------------------------------------------------------------------------------------------ intermediates
    def intermediates(self) -> dict:
        """Copies of every stage of the LAST forward (synchronises; tests / debugging)."""
        v, (N, B) = self._view, self._sizes
        dev = self._ws.device

        def arr(p, n, dtype):
            if n == 0:
                return torch.empty(0, dtype=dtype, device=dev)
            size = n * (4 if dtype in (torch.float32, torch.int32) else 8)
            off = p - self._ws.data_ptr()
            return self._ws[off:off + size].view(dtype).clone()
        rowptr1 = arr(v.rowptr1, N + 1, torch.int32)
        E1 = int(rowptr1[-1])
        ptr1p = arr(v.ptr1p, B + 1, torch.int32)
        M = int(ptr1p[-1])
        rowptr2 = arr(v.rowptr2, N + 1, torch.int32)[:M + 1]
        E2 = int(rowptr2[-1])
        return {
            "x0": arr(v.x0, N * 64, torch.float32).view(N, 64), "rowptr1": rowptr1,
            "col1": arr(v.col1, E1, torch.int32), "row1": arr(v.row1, E1, torch.int32),
            "x1": arr(v.x1, N * 64, torch.float32).view(N, 64), "weight1": arr(v.weight1, E1, torch.float32),
            "cluster1": arr(v.cluster1, N, torch.int32), "partner1": arr(v.partner1, N, torch.int32),
            "x1p": arr(v.x1p, M * 64, torch.float32).view(M, 64), "ptr1p": ptr1p, "rowptr2": rowptr2,
            "col2": arr(v.col2, E2, torch.int32), "row2": arr(v.row2, E2, torch.int32),
            "x2": arr(v.x2, M * 64, torch.float32).view(M, 64), "pooled": arr(v.pooled, B * 64, torch.float32).view(B, 64),
            "asel1": arr(v.asel1, N * 64, torch.int32).view(N, 64), "asel2": arr(v.asel2, M * 64, torch.int32).view(M, 64),
        }


DRN_NORM = [1. / 2950.0, 1. / 2950.0, 1. / 2950.0, 1. / 5.265625, 1. / 143.875, 1. / 589., 1. / 1.2050781, 1. / 211., 1.,
            1. / 7., 1.]


class DRNNet(nn.Module):
    """model/net.py:12-37 (commented out in the reference): the DRN regressing (MET, MET phi)."""

    def __init__(self):
        super().__init__()
        self.drn = DynamicReductionNetwork(input_dim=11, hidden_dim=64, k=8, output_dim=2, aggr="max",
                                           norm=torch.tensor(DRN_NORM))

    def forward(self, data, max_event_nodes=None):
        if self.drn._wants_grad() and self.drn.training and self.drn.edge_precision == "fp32":   # training: the wrapper's activations as (differentiable) torch ops on [B,2]
            import math
            output = self.drn(data, max_event_nodes)
            met = nn.functional.softplus(output[:, 0]).unsqueeze(1)
            metphi = math.pi * (2 * torch.sigmoid(output[:, 1]) - 1).unsqueeze(1)
            return torch.cat((met, metphi), 1)
        x, p, men = self.drn._prep(data, max_event_nodes)
        _, net_out = self.drn._launch(x, p, men, want_net_out=True)
        return net_out
