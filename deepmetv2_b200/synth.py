"""Seeded synthetic DY/TT-shaped and high-pile-up PF-candidate events.

The CMS samples the reference trains on are fetched over xrootd
(/root/reference/data_dytt/filelist.json) and are not available, so the bench and the
parity tests run on synthetic events with the reference's column layout
(/root/reference/model/data_loader.py:70-77):

    x[:, 0:8]  = pX, pY, pT, eta, d0, dz, mass, puppiWeight        (continuous)
    x[:, 8:11] = pdgId, charge, fromPV                              (categorical, stored as f32)
    y[B, 11]   = genMET px,py | PF MET px,py | PUPPI px,py | DeepMET resp px,py | reso px,py | HT

Host-side numpy only (np.random.default_rng => identical events on every machine).
"""
from __future__ import annotations

import dataclasses

import numpy as np
import torch

# name -> (mean n, sigma n, n_min, n_max, (jets_lo, jets_hi))
SHAPES = {
    "dytt": (1700.0, 250.0, 800, 3000, (2, 6)),
    "highpu": (5000.0, 500.0, 3000, 7000, (4, 6)),
    "tiny": (40.0, 10.0, 8, 80, (1, 2)),
    # Z -> nu nu (BASELINE config 4, the DRN variant): DY-like pile-up multiplicity, fewer hard jets (recoil only)
    "znunu": (1700.0, 250.0, 800, 3000, (1, 3)),
}

_PDG_VALUES = np.array([211, -211, 130, 22, 11, -11, 13, -13, 1, 2], dtype=np.float32)
_PDG_PROBS = np.array([0.30, 0.30, 0.10, 0.20, 0.01, 0.01, 0.005, 0.005, 0.04, 0.03])
_PDG_PROBS = _PDG_PROBS / _PDG_PROBS.sum()
_MASS = {211: 0.13957, 130: 0.49761, 22: 0.0, 11: 0.000511, 13: 0.10566, 1: 0.0, 2: 0.0}


@dataclasses.dataclass
class EventBatch:
    """A collated mini-batch in the reference's PyG ``Batch`` field names (data.x/.y/.batch)."""
    x: torch.Tensor        # [N, 11] f32
    y: torch.Tensor        # [B, 11] f32
    batch: torch.Tensor    # [N] i64, sorted
    ptr: torch.Tensor      # [B+1] i32

    @property
    def num_graphs(self) -> int:
        return int(self.ptr.numel() - 1)

    def to(self, device, non_blocking: bool = False) -> "EventBatch":
        return EventBatch(*(t.to(device, non_blocking=non_blocking)
                            for t in (self.x, self.y, self.batch, self.ptr)))

    def pin_memory(self) -> "EventBatch":
        return EventBatch(*(t.pin_memory() for t in (self.x, self.y, self.batch, self.ptr)))


def _wrap_phi(phi):
    return (phi + np.pi) % (2.0 * np.pi) - np.pi


def make_event(rng: np.random.Generator, shape: str = "dytt", n: int | None = None) -> np.ndarray:
    mu, sigma, lo, hi, (jl, jh) = SHAPES[shape]
    if n is None:
        n = int(np.clip(np.rint(rng.normal(mu, sigma)), lo, hi))
    n_hf = int(0.15 * n)
    n_jet = int(0.15 * n)
    n_pu = n - n_hf - n_jet
    eta = np.empty(n)
    phi = np.empty(n)
    eta[:n_pu] = rng.uniform(-3.0, 3.0, n_pu)
    phi[:n_pu] = rng.uniform(-np.pi, np.pi, n_pu)
    eta[n_pu:n_pu + n_hf] = rng.uniform(3.0, 5.0, n_hf) * rng.choice([-1.0, 1.0], n_hf)
    phi[n_pu:n_pu + n_hf] = rng.uniform(-np.pi, np.pi, n_hf)
    n_jets = int(rng.integers(jl, jh + 1))
    n_jets = max(n_jets, 1)
    c_eta = rng.uniform(-2.0, 2.0, n_jets)
    c_phi = rng.uniform(-np.pi, np.pi, n_jets)
    which = rng.integers(0, n_jets, n_jet)
    eta[n_pu + n_hf:] = c_eta[which] + rng.normal(0.0, 0.15, n_jet)
    phi[n_pu + n_hf:] = _wrap_phi(c_phi[which] + rng.normal(0.0, 0.15, n_jet))

    pt = 0.5 + rng.exponential(1.5, n)
    pdg = rng.choice(_PDG_VALUES, size=n, p=_PDG_PROBS)
    pdg[n_pu:n_pu + n_hf] = rng.choice(np.array([1.0, 2.0], dtype=np.float32), n_hf)
    apdg = np.abs(pdg).astype(np.int64)
    charge = np.where(np.isin(apdg, (211, 11, 13)), np.sign(pdg) * np.where(apdg == 211, 1.0, -1.0), 0.0)
    mass = np.zeros(n)
    for k, v in _MASS.items():
        mass[apdg == k] = v
    puppi = np.where(charge != 0, rng.integers(0, 2, n).astype(np.float64), rng.uniform(0.0, 1.0, n))
    d0 = np.clip(rng.normal(0.0, 0.1, n), -5000.0, 5000.0)
    dz = np.clip(rng.normal(0.0, 0.5, n), -5000.0, 5000.0)
    from_pv = rng.integers(0, 4, n).astype(np.float64)

    x = np.stack([pt * np.cos(phi), pt * np.sin(phi), pt, eta, d0, dz, mass, puppi,
                  pdg.astype(np.float64), charge, from_pv], axis=1).astype(np.float32)
    return x[rng.permutation(n)]


def make_batch(num_events: int, shape: str = "dytt", seed: int = 0,
               sizes: list[int] | None = None) -> EventBatch:
    rng = np.random.default_rng(seed)
    xs = [make_event(rng, shape, None if sizes is None else sizes[i]) for i in range(num_events)]
    counts = np.array([len(a) for a in xs], dtype=np.int64)
    x = np.concatenate(xs, axis=0) if xs else np.zeros((0, 11), np.float32)
    y = rng.normal(0.0, 30.0, (num_events, 11)).astype(np.float32)
    y[:, 10] = np.abs(y[:, 10]) * 10.0  # HT
    batch = np.repeat(np.arange(num_events, dtype=np.int64), counts)
    ptr = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
    return EventBatch(torch.from_numpy(x), torch.from_numpy(y), torch.from_numpy(batch),
                      torch.from_numpy(ptr))


def boundary_ambiguous_pairs(etaphi: torch.Tensor, ptr: torch.Tensor, r: float = 0.4, ulps: int = 4) -> int:
    """Number of in-event pairs whose fp32 d^2 lies within ``ulps`` of r^2 — pairs for which the
    upstream FMA-vs-non-FMA choice could flip the edge (SURVEY.md §7 'margin check')."""
    r2 = np.float32(float(r) * float(r))
    tol = np.spacing(r2) * ulps
    pos = etaphi.cpu().numpy().astype(np.float32)
    p = ptr.cpu().numpy()
    cnt = 0
    for e in range(len(p) - 1):
        q = pos[p[e]:p[e + 1]]
        d = q[:, None, :] - q[None, :, :]
        d2 = (d * d).sum(-1, dtype=np.float32)
        cnt += int((np.abs(d2 - r2) <= tol).sum())
    return cnt
