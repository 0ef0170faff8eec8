"""Defines the neural network, loss function and metrics — the reference's ``model/net.py`` surface
(/root/reference/model/net.py:38-47 ``Net``, :49-62 ``loss_fn``, :71-90 ``u_perp_par_loss``, :92-157
``resolution``, :160-162 ``metrics``) on top of libgraphmet.so.  Same names, arguments and return types
(0-dim loss tensor, numpy arrays from ``resolution``)."""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn

from .. import _lib, ops
from .._lib import check, lib, ptr, stream
from .graph_met_network import GraphMETNetwork


class Net(nn.Module):
    def __init__(self, continuous_dim, categorical_dim):
        super(Net, self).__init__()
        self.graphnet = GraphMETNetwork(continuous_dim, categorical_dim,
                                        output_dim=1, hidden_dim=32,
                                        conv_depth=2)

    def forward(self, x_cont, x_cat, edge_index, batch):
        # torch.sigmoid(self.graphnet(...)) of the reference, with the sigmoid fused into the head kernel
        return self.graphnet._run(x_cont, x_cat, edge_index, batch, apply_sigmoid=True)


def _ptr_of(batch, num_events):
    """int32 event boundaries from the reference's ``batch`` argument (sorted int64 [N]); a ptr tensor
    attached by our own loader (``batch._gm_ptr``) is used as is."""
    p = getattr(batch, "_gm_ptr", None)
    if p is not None and p.numel() == num_events + 1 and p.device == batch.device:
        return p
    return ops.batch_to_ptr(batch, num_events)


class _MetLoss(torch.autograd.Function):
    """loss_fn as one node: per-event segmented MET reduction + loss, gradient produced in the forward."""

    @staticmethod
    def forward(ctx, weights, prediction, truth, ptr_i32):
        N, B = weights.shape[0], truth.shape[0]
        dev = weights.device
        met = torch.empty(B, 2, dtype=torch.float32, device=dev)
        loss = torch.empty((), dtype=torch.float32, device=dev)
        gw = torch.empty(N, dtype=torch.float32, device=dev) if ctx.needs_input_grad[0] else None
        check(lib().gm_met_loss(ptr(weights), ptr(prediction), prediction.stride(0), ptr(truth), truth.stride(0),
                                ptr(ptr_i32), N, B, ptr(met), ptr(loss), ptr(gw), stream()), "gm_met_loss")
        ctx.gw = gw
        ctx.met = met
        return loss

    @staticmethod
    def backward(ctx, grad_loss):
        return ctx.gw * grad_loss, None, None, None


def _check_loss_args(weights, prediction, truth):
    _lib.require_cuda(weights, "weights")
    if prediction.dtype != torch.float32 or prediction.stride(1) != 1:
        prediction = prediction.float().contiguous()
    if truth.dtype != torch.float32 or truth.stride(1) != 1:
        truth = truth.float().contiguous()
    return weights.contiguous().float(), prediction, truth


def loss_fn(weights, prediction, truth, batch):
    """0.5 * mean_events((METx + true_px)^2 + (METy + true_py)^2), MET = sum_n w_n * (px, py)."""
    weights, prediction, truth = _check_loss_args(weights, prediction, truth)
    return _MetLoss.apply(weights, prediction, truth, _ptr_of(batch, truth.shape[0]))


def met_from_weights(weights, prediction, batch, num_events):
    """[B,2] per-event (sum w*px, sum w*py): the two scatter_add calls of model/net.py:55-56."""
    ptr_i32 = _ptr_of(batch, num_events)
    w = weights.detach().contiguous().float()
    metx = ops.segment_sum(w * prediction[:, 0], ptr_i32)
    mety = ops.segment_sum(w * prediction[:, 1], ptr_i32)
    return torch.stack((metx, mety), dim=1)


def u_perp_par_loss(weights, prediction, truth, batch):
    """model/net.py:71-90, including its qTy = truth[:,0] (sic). Unused by train.py; torch glue over the
    segmented MET kernel (not differentiable w.r.t. weights)."""
    weights, prediction, truth = _check_loss_args(weights, prediction, truth)
    met = met_from_weights(weights, prediction, batch, truth.shape[0])
    v_qT = torch.stack((truth[:, 0], truth[:, 0]), dim=1)
    vector = -met
    response = (vector * v_qT).sum(1) / (v_qT * v_qT).sum(1)
    v_paral = response[:, None] * v_qT
    u_paral = v_paral.norm(dim=1) - v_qT.norm(dim=1)
    u_perp = (vector - v_paral).norm(dim=1)
    return 0.5 * (u_paral ** 2 + u_perp ** 2).mean()


_RES_NAMES = ["MET", "pfMET", "puppiMET", "deepMETResponse", "deepMETResolution"]


def resolution(weights, prediction, truth, batch):
    """-> ({name: [u_perp, u_par, response] (numpy [B] each)}, qT numpy [B]) like model/net.py:92-157,
    computed by two kernels and ONE device->host copy instead of 10-16."""
    weights, prediction, truth = _check_loss_args(weights.detach(), prediction, truth)
    B = truth.shape[0]
    ptr_i32 = _ptr_of(batch, B)
    dev = weights.device
    met = torch.empty(B, 2, dtype=torch.float32, device=dev)
    scratch = torch.empty((), dtype=torch.float32, device=dev)
    l = lib()
    check(l.gm_met_loss(ptr(weights), ptr(prediction), prediction.stride(0), ptr(truth), truth.stride(0),
                        ptr(ptr_i32), weights.shape[0], B, ptr(met), ptr(scratch), None, stream()), "gm_met_loss")
    K = 5 if truth.size()[1] > 6 else 3
    out = torch.empty(K * 3 + 1, B, dtype=torch.float32, device=dev)
    check(l.gm_resolution(ptr(met), ptr(truth), truth.stride(0), truth.shape[1], B, ptr(out), ptr(out[K * 3]),
                          stream()), "gm_resolution")
    host = out.cpu().numpy()
    resolutions = {_RES_NAMES[k]: [host[3 * k], host[3 * k + 1], host[3 * k + 2]] for k in range(K)}
    return resolutions, host[K * 3]


# maintain all metrics required in this dictionary- these are used in the training and evaluation loops
metrics = {
    'resolution': resolution,
}
