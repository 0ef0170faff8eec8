"""``DynamicReductionNetwork`` with the reference's constructor / forward signature and state_dict layout
(/root/reference/model/dynamic_reduction_network.py:32-104) and the ``Net`` wrapper the reference keeps commented out at
/root/reference/model/net.py:12-37 (``DRNNet`` here), executed by libgraphmet.so (``gm_drn_forward``, csrc/drn_net.cu +
csrc/drn_graph.cu + csrc/knn.cu).

The module tree (``datanorm``, ``inputnet.{0,2,4}``, ``edgeconv{1,2}.nn.{0,2,4}``, ``output.{0,2,4}``) is the
reference's, so a checkpoint of the reference model strict-loads.  ``forward(data)`` takes any object with ``.x`` [N,11]
and ``.batch`` [N] (sorted, PyG collation) or ``.ptr`` [B+1]; it does not execute the sub-modules but hands their
storages to ONE C call that runs the whole forward without a host synchronisation.  Forward only (inference and the
train-mode BatchNorm statistics); there is no CPU fallback.
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
        self._view = None
        self._sizes = (0, 0)

    # ------------------------------------------------------------------------------------------ C structs
    def _param_struct(self) -> GmDrnParams:
        if self.input_dim != 11 or self.hidden_dim != 64 or self.edgeconv1.aggr != "max" or self.output_dim > 8:
            raise NotImplementedError("the sm_100a DRN kernels are specialised for the configuration of "
                                      "model/net.py:14-29: input_dim=11, hidden_dim=64, aggr='max'")
        P = GmDrnParams()
        P.datanorm = ptr(self.datanorm)
        for i, idx in enumerate((0, 2, 4)):
            P.in_w[i], P.in_b[i] = ptr(self.inputnet[idx].weight), ptr(self.inputnet[idx].bias)
            P.out_w[i], P.out_b[i] = ptr(self.output[idx].weight), ptr(self.output[idx].bias)
        for i, ec in enumerate((self.edgeconv1, self.edgeconv2)):
            c, seq = GmDrnConv(), ec.nn
            c.w1, c.b1, c.w2, c.b2 = ptr(seq[0].weight), ptr(seq[0].bias), ptr(seq[2].weight), ptr(seq[2].bias)
            bn = seq[4]
            c.bn_weight, c.bn_bias = ptr(bn.weight), ptr(bn.bias)
            c.bn_running_mean, c.bn_running_var = ptr(bn.running_mean), ptr(bn.running_var)
            c.bn_num_batches_tracked = ptr(bn.num_batches_tracked)
            P.conv[i] = c
        P.output_dim, P.k = int(self.output_dim), int(self.k)
        return P

    # ------------------------------------------------------------------------------------------ forward
    def _run(self, data, want_net_out: bool, max_event_nodes=None):
        x = data.x
        _lib.require_cuda(x, "data.x")
        if x.dtype != torch.float32 or x.dim() != 2 or x.shape[1] < 11 or x.stride(1) != 1:
            raise ValueError("data.x must be float32 [N, >=11] with unit column stride")
        p = getattr(data, "ptr", None)
        if p is None or p.dtype != torch.int32:
            batch = data.batch
            B = int(batch[-1].item()) + 1
            p = ops.batch_to_ptr(batch, B)
        B, N = p.numel() - 1, x.shape[0]
        if max_event_nodes is None:
            max_event_nodes = getattr(data, "max_event_nodes", None)
        if max_event_nodes is None:
            max_event_nodes = int((p[1:] - p[:-1]).max().item())
        if torch.is_grad_enabled() and any(q.requires_grad for q in self.parameters()) and self.training:
            pass    # forward statistics only: the DRN backward is not part of this library
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
        return out, net_out

    def forward(self, data, max_event_nodes=None):
        out, _ = self._run(data, False, max_event_nodes)
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
        _, net_out = self.drn._run(data, True, max_event_nodes)
        return net_out
