"""Packed ragged event batches and their double-buffered host -> device feed.

The reference loads ONE ``.pt`` file per event and lets PyG's ``DataLoader`` collate them on the host
(``model/data_loader.py:59-61,107-110``), then ``data.to(device)`` inside the step (``train.py:41``).  Here a batch is
three flat arrays -- ``x[N,11] f32``, ``y[B,11] f32``, ``ptr[B+1] i32`` -- packed once (:func:`pack_events`), stored as a
shard (:func:`save_shard` / :func:`load_shard`), pinned, and copied by :class:`HostFeeder` on a dedicated copy
stream into a ring of device buffers while the previous batch is still being computed on; the consumer sees
``EventBatch`` objects with the reference's field names (``.x / .y / .batch / .ptr``).
"""
from __future__ import annotations

from collections import deque
from typing import Iterable, Iterator, Sequence

import numpy as np
import torch

from .synth import EventBatch


def pack_events(xs: Sequence[torch.Tensor], ys: Sequence[torch.Tensor], pin: bool = True) -> EventBatch:
    """Collate per-event ``x_i [n_i, 11]`` / ``y_i [11]`` (what ``METDataset.get`` returns per file) into one packed
    batch; ``batch`` (PyG's node -> event vector) is derived from ``ptr``."""
    if len(xs) != len(ys):
        raise ValueError("one truth row per event")
    counts = torch.tensor([int(x.shape[0]) for x in xs], dtype=torch.int64)
    ptr = torch.zeros(len(xs) + 1, dtype=torch.int32)
    ptr[1:] = torch.cumsum(counts, 0).to(torch.int32)
    x = torch.cat([x.reshape(-1, x.shape[-1]).float() for x in xs], 0) if len(xs) else torch.zeros(0, 11)
    y = torch.stack([y.reshape(-1).float() for y in ys], 0) if len(ys) else torch.zeros(0, 11)
    batch = torch.repeat_interleave(torch.arange(len(xs), dtype=torch.int64), counts)
    b = EventBatch(x.contiguous(), y.contiguous(), batch, ptr)
    return b.pin_memory() if pin and torch.cuda.is_available() else b


def save_shard(path: str, b: EventBatch) -> None:
    """One file per BATCH (or per many batches' worth of events) instead of one per event."""
    np.savez(path, x=b.x.numpy(), y=b.y.numpy(), ptr=b.ptr.numpy())


def load_shard(path: str, pin: bool = True) -> EventBatch:
    with np.load(path) as z:
        x, y, ptr = torch.from_numpy(z["x"]), torch.from_numpy(z["y"]), torch.from_numpy(z["ptr"]).int()
    counts = (ptr[1:] - ptr[:-1]).long()
    b = EventBatch(x, y, torch.repeat_interleave(torch.arange(counts.numel(), dtype=torch.int64), counts), ptr)
    return b.pin_memory() if pin and torch.cuda.is_available() else b


class _Slot:
    def __init__(self):
        self.x = self.y = self.ptr = None
        self.ready = torch.cuda.Event()        # copies of this slot have landed
        self.consumed = torch.cuda.Event()     # the compute stream is done with this slot
        self.used = False


class HostFeeder:
    """Iterate device-resident batches over an iterable of (pinned) host batches, copying batch k+1 on a copy
    stream while batch k is computed on.  ``depth`` device buffers are kept and re-used (grown on demand), so the
    steady state allocates nothing.  Yields ``EventBatch(x, y, None, ptr)`` with ``.max_event_nodes`` set from the
    HOST ptr (no device sync); the views are valid until the next ``next()`` call + ``depth - 1`` steps.

    The consumer must enqueue its work on the stream that is current when it calls ``next()``."""

    def __init__(self, batches: Iterable[EventBatch], device, depth: int = 2):
        if depth < 2:
            raise ValueError("depth >= 2 (one buffer being computed on, one being filled)")
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("HostFeeder feeds a CUDA device (no CPU fallback)")
        self.batches = batches
        self.copy_stream = torch.cuda.Stream(self.device)
        self.slots = [_Slot() for _ in range(depth)]

    @staticmethod
    def _fit(buf, like: torch.Tensor, device):
        n = like.shape[0]
        if buf is None or buf.shape[0] < n or buf.shape[1:] != like.shape[1:]:
            buf = torch.empty((max(n, 1),) + tuple(like.shape[1:]), dtype=like.dtype, device=device)
        return buf

    def _prefetch(self, slot: _Slot, hb: EventBatch):
        with torch.cuda.stream(self.copy_stream):
            if slot.used:
                self.copy_stream.wait_event(slot.consumed)
            slot.x = self._fit(slot.x, hb.x, self.device)
            slot.y = self._fit(slot.y, hb.y, self.device)
            slot.ptr = self._fit(slot.ptr, hb.ptr, self.device)
            N, B = hb.x.shape[0], hb.y.shape[0]
            slot.x[:N].copy_(hb.x, non_blocking=True)
            slot.y[:B].copy_(hb.y, non_blocking=True)
            slot.ptr[:B + 1].copy_(hb.ptr, non_blocking=True)
            slot.ready.record(self.copy_stream)
        slot.used = True
        men = getattr(hb, "max_event_nodes", None)                   # host tensor: no device sync either way
        if men is None:
            men = int((hb.ptr[1:] - hb.ptr[:-1]).max()) if B else 0
        return N, B, men

    def __iter__(self) -> Iterator[EventBatch]:
        it = iter(self.batches)
        inflight: deque = deque()
        free = deque(self.slots)
        while free:
            hb = next(it, None)
            if hb is None:
                break
            s = free.popleft()
            inflight.append((s, self._prefetch(s, hb)))
        while inflight:
            s, (N, B, men) = inflight.popleft()
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(s.ready)
            out = EventBatch(s.x[:N], s.y[:B], None, s.ptr[:B + 1])
            out.max_event_nodes = men
            yield out
            s.consumed.record(torch.cuda.current_stream(self.device))
            hb = next(it, None)
            if hb is not None:
                inflight.append((s, self._prefetch(s, hb)))
