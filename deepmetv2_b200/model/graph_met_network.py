"""``GraphMETNetwork`` with the reference's constructor / forward signature and state_dict layout
(/root/reference/model/graph_met_network.py:11-69), executed by libgraphmet.so.

The module tree (``embed_charge``, ``embed_continuous.0``, ``conv_continuous.{l}.0.nn.0``, ``bn_all`` …) is
the reference's, so its checkpoints strict-load and ``torch.optim.AdamW(model.parameters())`` sees the same
23 tensors in the same order.  ``forward`` does not execute those sub-modules: it hands their storages to
one fused CUDA forward (``gm_net_forward``) and registers one hand-written backward (``gm_net_backward``).
"""
from __future__ import annotations

import ctypes as C

import torch
from torch import nn

from .. import _lib, ops
from .._lib import AGGR, PRECISION, GmBatchNorm, GmParams, check, lib, ptr, stream


class EdgeConv(nn.Module):
    """Holder with torch_geometric.nn.EdgeConv's interface: ``nn`` (the edge MLP, here one Linear(2H, H)),
    ``aggr`` (default 'max' like upstream) and ``jittable()``.  Standalone ``forward`` runs the CUDA
    EdgeConv kernel (inference only); inside GraphMETNetwork the layer is part of the fused pass."""

    def __init__(self, nn, aggr="max"):  # noqa: A002  (upstream argument name)
        super().__init__()
        self.nn = nn
        self.aggr = aggr
        self.flow = "source_to_target"
        self.precision = "fp32"

    def jittable(self):
        return self

    def forward(self, x, edge_index):
        _lib.require_cuda(x, "x")
        lin = self.nn[0]
        if len(self.nn) != 1 or lin.in_features != 2 * lin.out_features or lin.out_features != 32:
            raise NotImplementedError("EdgeConv kernels are specialised for nn = Sequential(Linear(64, 32))")
        if torch.is_grad_enabled() and (x.requires_grad or lin.weight.requires_grad):
            raise NotImplementedError("standalone EdgeConv is forward-only; train through GraphMETNetwork")
        csr = edge_index if isinstance(edge_index, ops.GraphCSR) else ops.csr_from_edge_index(edge_index, x.shape[0])
        x = x.contiguous().float()
        out = torch.empty_like(x)
        l = lib()
        ws = torch.empty(l.gm_edgeconv_ws_bytes(x.shape[0]), dtype=torch.uint8, device=x.device)
        check(l.gm_edgeconv_forward(ptr(x), ptr(csr.rowptr), ptr(csr.col), x.shape[0], ptr(lin.weight),
                                    ptr(lin.bias), AGGR[self.aggr], PRECISION[self.precision], ptr(ws), ws.numel(),
                                    ptr(out), None, stream()), "gm_edgeconv_forward")
        return out


def _bn_struct(bn: nn.BatchNorm1d | None, grads: dict | None = None, prefix: str = "") -> GmBatchNorm:
    s = GmBatchNorm()
    if grads is not None:
        s.weight, s.bias = ptr(grads[prefix + "weight"]), ptr(grads[prefix + "bias"])
        return s
    s.weight, s.bias = ptr(bn.weight), ptr(bn.bias)
    s.running_mean, s.running_var = ptr(bn.running_mean), ptr(bn.running_var)
    s.num_batches_tracked = ptr(bn.num_batches_tracked)
    return s


class _GraphMETFunction(torch.autograd.Function):
    """One autograd node for the whole network: 11 forward launches, 16 backward launches."""

    @staticmethod
    def forward(ctx, net, x_cont, x_cat, csr, apply_sigmoid, save, *params):
        N = x_cont.shape[0]
        dev = x_cont.device
        depth = len(net.conv_continuous)
        l = lib()
        ws = torch.empty(l.gm_net_ws_bytes(N, depth, int(save)), dtype=torch.uint8, device=dev)
        out = torch.empty(N, dtype=torch.float32, device=dev)
        P = net._param_struct()
        cat_i64 = x_cat if x_cat.dtype == torch.int64 else None
        cat_f32 = x_cat if x_cat.dtype == torch.float32 else None
        training = bool(net.training)
        check(l.gm_net_forward(C.byref(P), ptr(x_cont), x_cont.stride(0), ptr(cat_i64), ptr(cat_f32),
                               0 if cat_f32 is None else cat_f32.stride(0), ptr(csr.rowptr), ptr(csr.col), N,
                               None, int(training), int(save), int(apply_sigmoid), AGGR[net.aggr],
                               PRECISION[net.edge_precision], float(net.bn_all.momentum), float(net.bn_all.eps),
                               ptr(ws), ws.numel(), ptr(out), ptr(net._err_flags(dev)), stream()), "gm_net_forward")
        ctx.net, ctx.csr, ctx.ws, ctx.saved_save = net, csr, ws, save
        ctx.apply_sigmoid, ctx.training = bool(apply_sigmoid), training
        ctx.x_cont, ctx.x_cat, ctx.out = x_cont, x_cat, out
        return out

    @staticmethod
    def backward(ctx, grad_out):
        net, csr = ctx.net, ctx.csr
        if not ctx.saved_save:
            raise RuntimeError("GraphMETNetwork forward ran without save_for_backward")
        N = ctx.x_cont.shape[0]
        dev = ctx.x_cont.device
        depth = len(net.conv_continuous)
        l = lib()
        params = list(net.parameters())
        flat = torch.empty(sum(p.numel() for p in params), dtype=torch.float32, device=dev)   # every gradient is overwritten
        views, off = [], 0
        for p in params:
            views.append(flat[off:off + p.numel()].view_as(p))
            off += p.numel()
        G = net._param_struct(grad_views=views)
        P = net._param_struct()
        bws = torch.empty(l.gm_net_bwd_ws_bytes(N, depth), dtype=torch.uint8, device=dev)
        rowptr_t, col_t = csr.transposed()
        x_cat = ctx.x_cat
        cat_i64 = x_cat if x_cat.dtype == torch.int64 else None
        cat_f32 = x_cat if x_cat.dtype == torch.float32 else None
        grad_out = grad_out.contiguous().float()
        check(l.gm_net_backward(C.byref(P), C.byref(G), ptr(ctx.x_cont), ctx.x_cont.stride(0), ptr(cat_i64),
                                ptr(cat_f32), 0 if cat_f32 is None else cat_f32.stride(0), ptr(csr.rowptr),
                                ptr(csr.col), ptr(rowptr_t), ptr(col_t), ptr(csr.capped_flag), N, None, int(ctx.training),
                                int(ctx.apply_sigmoid), AGGR[net.aggr], float(net.bn_all.eps), ptr(ctx.ws),
                                ctx.ws.numel(), ptr(bws), bws.numel(), ptr(ctx.out), ptr(grad_out), stream()),
              "gm_net_backward")
        grads = tuple(v if p.requires_grad else None for v, p in zip(views, params))
        return (None, None, None, None, None, None) + grads


class GraphMETNetwork(nn.Module):
    def __init__(self, continuous_dim, cat_dim, output_dim=1, hidden_dim=32, conv_depth=1):
        super().__init__()
        if hidden_dim != 32 or output_dim != 1:
            raise NotImplementedError("the sm_100a kernels are specialised for hidden_dim=32, output_dim=1 "
                                      "(model/net.py:41-43)")
        if not 1 <= continuous_dim <= 8 or conv_depth > _lib.MAX_CONV_DEPTH:
            raise NotImplementedError("continuous_dim must be in 1..8 and conv_depth <= 4")
        self.continuous_dim, self.cat_dim = continuous_dim, cat_dim

        self.embed_charge = nn.Embedding(3, hidden_dim // 4)
        self.embed_pdgid = nn.Embedding(7, hidden_dim // 4)
        self.embed_pv = nn.Embedding(8, hidden_dim // 4)
        self.embed_continuous = nn.Sequential(nn.Linear(continuous_dim, hidden_dim // 2), nn.ELU())
        self.embed_categorical = nn.Sequential(nn.Linear(3 * hidden_dim // 4, hidden_dim // 2), nn.ELU())
        self.encode_all = nn.Sequential(nn.Linear(hidden_dim, hidden_dim), nn.ELU())
        self.bn_all = nn.BatchNorm1d(hidden_dim)
        self.conv_continuous = nn.ModuleList()
        for _ in range(conv_depth):
            mesg = nn.Sequential(nn.Linear(2 * hidden_dim, hidden_dim))
            self.conv_continuous.append(nn.ModuleList())
            self.conv_continuous[-1].append(EdgeConv(nn=mesg).jittable())
            self.conv_continuous[-1].append(nn.BatchNorm1d(hidden_dim))
        self.output = nn.Sequential(nn.Linear(hidden_dim, hidden_dim // 2), nn.ELU(),
                                    nn.Linear(hidden_dim // 2, output_dim))
        self.pdgs = [1, 2, 11, 13, 22, 130, 211]
        self.edge_precision = "fp32"     # 'fp32' (parity) | 'tf32' | 'bf16' | 'tf32x3'  (tcgen05 variants)
        self._flags = None
        self._flags_host = None
        self._flags_event = None

    # ------------------------------------------------------------------ plumbing
    @property
    def aggr(self) -> str:
        return self.conv_continuous[0][0].aggr if len(self.conv_continuous) else "max"

    def _err_flags(self, dev):
        if self._flags is None or self._flags.device != dev:
            self._flags = torch.zeros(1, dtype=torch.int32, device=dev)
            self._flags_host = torch.zeros(1, dtype=torch.int32).pin_memory()
            self._flags_event = None
        return self._flags

    def check_errors(self, wait: bool = True) -> None:
        """Raise IndexError if a categorical index left its embedding table (the reference raises from
        nn.Embedding, model/graph_met_network.py:49-55).  Reported asynchronously, like a CUDA assert."""
        if self._flags is None:
            return
        if wait:
            flags = int(self._flags.item())
        else:
            if self._flags_event is None or not self._flags_event.query():
                return
            flags = int(self._flags_host.item())
        if flags & _lib.FLAG_CAT_OUT_OF_RANGE:
            self._flags.zero_()
            raise IndexError("index out of range in self (embed_charge / embed_pdgid / embed_pv)")

    def _param_struct(self, grad_views=None) -> GmParams:
        P = GmParams()
        P.continuous_dim = self.continuous_dim
        P.conv_depth = len(self.conv_continuous)
        if grad_views is not None:
            it = iter(grad_views)
            nxt = lambda: ptr(next(it))  # noqa: E731  (parameters() order == reference construction order)
            P.embed_charge, P.embed_pdgid, P.embed_pv = nxt(), nxt(), nxt()
            P.cont_w, P.cont_b = nxt(), nxt()
            P.cat_w, P.cat_b = nxt(), nxt()
            P.enc_w, P.enc_b = nxt(), nxt()
            P.bn_all.weight, P.bn_all.bias = nxt(), nxt()
            for li in range(P.conv_depth):
                P.conv_w[li], P.conv_b[li] = nxt(), nxt()
                P.conv_bn[li].weight, P.conv_bn[li].bias = nxt(), nxt()
            P.out1_w, P.out1_b, P.out2_w, P.out2_b = nxt(), nxt(), nxt(), nxt()
            return P
        P.embed_charge, P.embed_pdgid, P.embed_pv = (ptr(self.embed_charge.weight), ptr(self.embed_pdgid.weight),
                                                      ptr(self.embed_pv.weight))
        P.cont_w, P.cont_b = ptr(self.embed_continuous[0].weight), ptr(self.embed_continuous[0].bias)
        P.cat_w, P.cat_b = ptr(self.embed_categorical[0].weight), ptr(self.embed_categorical[0].bias)
        P.enc_w, P.enc_b = ptr(self.encode_all[0].weight), ptr(self.encode_all[0].bias)
        P.bn_all = _bn_struct(self.bn_all)
        for li, co_conv in enumerate(self.conv_continuous):
            lin = co_conv[0].nn[0]
            P.conv_w[li], P.conv_b[li] = ptr(lin.weight), ptr(lin.bias)
            P.conv_bn[li] = _bn_struct(co_conv[1])
        P.out1_w, P.out1_b = ptr(self.output[0].weight), ptr(self.output[0].bias)
        P.out2_w, P.out2_b = ptr(self.output[2].weight), ptr(self.output[2].bias)
        return P

    def _run(self, x_cont, x_cat, edge_index, batch, apply_sigmoid):
        _lib.require_cuda(x_cont, "x_cont")
        for p in self.parameters():
            if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                raise _lib.GraphMETLibraryError("GraphMETNetwork parameters must be contiguous float32 CUDA "
                                                "tensors (call .to('cuda'))")
        self.check_errors(wait=False)
        if x_cont.dtype != torch.float32 or x_cont.dim() != 2 or x_cont.shape[1] != self.continuous_dim:
            raise ValueError(f"x_cont must be float32 [N,{self.continuous_dim}]")
        if x_cont.stride(1) != 1:
            x_cont = x_cont.contiguous()
        if x_cat.dtype == torch.int64:
            x_cat = x_cat.contiguous()
        elif x_cat.dtype == torch.float32:
            if x_cat.stride(1) != 1:
                x_cat = x_cat.contiguous()
        else:
            x_cat = x_cat.long().contiguous()
        if x_cat.shape != (x_cont.shape[0], 3):
            raise ValueError("x_cat must be [N,3] (pdgId, charge, fromPV)")
        csr = edge_index if isinstance(edge_index, ops.GraphCSR) else \
            ops.csr_from_edge_index(edge_index, x_cont.shape[0])
        params = list(self.parameters())
        save = bool(torch.is_grad_enabled() and any(p.requires_grad for p in params))
        out = _GraphMETFunction.apply(self, x_cont, x_cat, csr, apply_sigmoid, save, *params)
        if self._flags_host is not None:
            self._flags_host.copy_(self._flags, non_blocking=True)
            self._flags_event = torch.cuda.Event()
            self._flags_event.record()
        return out

    def forward(self, x_cont, x_cat, edge_index, batch):
        """-> Tensor[N] (pre-sigmoid), like the reference's ``out.squeeze(-1)``. ``batch`` is accepted and,
        as in the reference (model/graph_met_network.py:47-69), unused."""
        return self._run(x_cont, x_cat, edge_index, batch, apply_sigmoid=False)
