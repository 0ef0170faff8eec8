"""Evaluation-epoch metrics kept on the device: the accumulation loop of
/root/reference/evaluate.py:71-108 and the qT-binned summary of evaluate.py:110-156 without the reference's per-batch
device->host copies and host-side numpy concatenation / np.quantile.

    acc = ResolutionAccumulator(device)
    for data in loader:                                   # evaluate.py:71-108
        weights = model(...)
        acc.update(weights, data.x, data.y, data.batch)   # MET + projections appended in device memory, no sync
    resolution_hists = acc.summary()                      # evaluate.py:110-156: one launch, one small D2H

``summary()`` returns the reference's ``resolution_hists`` structure: ``{flavour: {'u_perp_resolution': (hist[40],
edges[41]), 'u_perp_scaled_resolution': ..., 'u_par_resolution': ..., 'u_par_scaled_resolution': ..., 'R': ...}}``.
"""
from __future__ import annotations

import numpy as np
import torch

from ._lib import check, lib, ptr, stream
from .model.net import _RES_NAMES, _check_loss_args, _ptr_of

_FIELDS = ["u_perp_resolution", "u_perp_scaled_resolution", "u_par_resolution", "u_par_scaled_resolution", "R"]


class ResolutionAccumulator:
    def __init__(self, device, capacity: int = 1 << 16, max_x: float = 400.0, bin_width: float = 10.0):
        self.device = torch.device(device)
        self.cap, self.count, self.K = int(capacity), 0, None
        self.max_x, self.bin_width = float(max_x), float(bin_width)
        self.vals = None          # [K*3 + 1, cap]: u_perp / u_par / R per flavour, then qT
        self.losses = []

    def _grow(self, need: int):
        cap = self.cap
        while cap < need:
            cap *= 2
        if self.vals is None or cap != self.vals.shape[1]:
            new = torch.empty(self.K * 3 + 1, cap, dtype=torch.float32, device=self.device)
            if self.vals is not None and self.count:
                new[:, :self.count] = self.vals[:, :self.count]
            self.vals, self.cap = new, cap

    def update(self, weights, prediction, truth, batch):
        """model/net.py:92-157 for one batch, appended to the epoch's device arrays (no host synchronisation)."""
        weights, prediction, truth = _check_loss_args(weights.detach(), prediction, truth)
        B = truth.shape[0]
        K = 5 if truth.shape[1] > 6 else 3
        if self.K is None:
            self.K = K
        elif self.K != K:
            raise ValueError("truth column count changed inside an epoch")
        self._grow(self.count + B)
        ptr_i32 = _ptr_of(batch, B)
        l, dev = lib(), self.device
        met = torch.empty(B, 2, dtype=torch.float32, device=dev)
        scratch = torch.empty((), dtype=torch.float32, device=dev)
        check(l.gm_met_loss(ptr(weights), ptr(prediction), prediction.stride(0), ptr(truth), truth.stride(0), ptr(ptr_i32),
                            weights.shape[0], B, ptr(met), ptr(scratch), None, stream()), "gm_met_loss")
        out = torch.empty(K * 3 + 1, B, dtype=torch.float32, device=dev)
        check(l.gm_resolution(ptr(met), ptr(truth), truth.stride(0), truth.shape[1], B, ptr(out), ptr(out[K * 3]), stream()),
              "gm_resolution")
        self.vals[:, self.count:self.count + B] = out
        self.count += B

    def summary(self) -> dict:
        """evaluate.py:110-156."""
        if not self.count:
            return {}
        n_bins = int(round(self.max_x / self.bin_width)) - 1          # np.arange(0, 400, 10): 40 edges, 39 bins
        out = torch.empty(self.K, 5, n_bins + 1, dtype=torch.float32, device=self.device)
        check(lib().gm_resolution_summary(ptr(self.vals), self.vals.stride(0), ptr(self.vals[self.K * 3]), self.count, self.K,
                                          self.bin_width, n_bins, ptr(out), stream()), "gm_resolution_summary")
        host = out.cpu().numpy()
        edges = np.linspace(0.0, self.max_x, n_bins + 2)
        return {_RES_NAMES[k]: {f: (host[k, i], edges) for i, f in enumerate(_FIELDS)} for k in range(self.K)}

    def arrays(self):
        """The epoch's per-event arrays on the host: ({flavour: [u_perp, u_par, R]}, qT) -- what the reference's loop
        builds with np.concatenate (one copy)."""
        host = self.vals[:, :self.count].cpu().numpy()
        return ({_RES_NAMES[k]: [host[3 * k], host[3 * k + 1], host[3 * k + 2]] for k in range(self.K)}, host[self.K * 3])
