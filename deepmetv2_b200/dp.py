"""Event-sharded data parallelism for the GraphMET training step.

The reference is single-device (train.py:72).  Events are independent in the graph build, EdgeConv and
the MET reduction, so the only place the path shards is the event axis: one process per GPU, every
rank owns a contiguous block of events, BatchNorm uses the rank-local batch statistics (plain DDP
semantics) and the ONE collective per step is a SUM all-reduce of the flat 6,641-float gradient
buffer, scaled by 1/world inside the fused AdamW kernel (``gm_adamw_step(grad_scale=1/world)``).

This module is host logic only (no kernels): the partitioner, the batch slicer and the collective
wrapper.  It works on any torch.distributed backend, which is how the CPU tests (gloo, world_size 2)
cover the N>1 path without a GPU.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from .synth import EventBatch


def event_cost(counts: torch.Tensor, edges_per_node: float = 0.0) -> torch.Tensor:
    """Work estimate per event: n (node-level kernels) + edges_per_node·n²/n̄ (the radius graph and the
    EdgeConv layers are O(E) and E grows ~ n² at fixed detector area)."""
    n = counts.to(torch.float64)
    if edges_per_node <= 0.0 or n.numel() == 0:
        return n
    return n + edges_per_node * n * n / n.mean().clamp_min(1.0)


def shard_events(counts, world: int, edges_per_node: float = 0.0) -> list[tuple[int, int]]:
    """Split events 0..B-1 into ``world`` CONTIGUOUS blocks [lo, hi) balanced by ``event_cost``.

    Contiguity keeps every rank's nodes one slice of ``x`` (no gather).  Greedy on the prefix sum:
    block k ends at the event whose cumulative cost is closest to (k+1)/world of the total.  Every
    rank gets at least one event when B >= world; with B < world the trailing ranks get empty blocks
    (they still join the all-reduce with a zero gradient)."""
    counts = torch.as_tensor(counts)
    B = int(counts.numel())
    if world < 1:
        raise ValueError("world must be >= 1")
    cost = event_cost(counts, edges_per_node)
    csum = torch.cat([cost.new_zeros(1), cost.cumsum(0)])
    total = float(csum[-1])
    bounds = [0]
    for k in range(1, world):
        target = total * k / world
        j = int(torch.searchsorted(csum, torch.tensor(target, dtype=csum.dtype)).item())
        if j > 0 and abs(float(csum[j - 1]) - target) <= abs(float(csum[min(j, B)]) - target):
            j -= 1
        lo_min = bounds[-1] + (1 if B >= world else 0)          # at least one event per rank if possible
        hi_max = B - (world - k) if B >= world else B
        bounds.append(max(lo_min, min(j, hi_max)))
    bounds.append(B)
    return [(bounds[k], bounds[k + 1]) for k in range(world)]


def slice_batch(batch: EventBatch, lo: int, hi: int) -> EventBatch:
    """Events [lo, hi) of a collated batch as a batch of its own (views; ptr/batch re-based)."""
    n0, n1 = int(batch.ptr[lo]), int(batch.ptr[hi])
    ptr = (batch.ptr[lo:hi + 1] - batch.ptr[lo]).contiguous()
    b = None if batch.batch is None else batch.batch[n0:n1] - lo
    return EventBatch(batch.x[n0:n1], batch.y[lo:hi], b, ptr)


def shard_batch(batch: EventBatch, rank: int, world: int, edges_per_node: float = 0.0) -> EventBatch:
    counts = (batch.ptr[1:] - batch.ptr[:-1]).cpu()
    lo, hi = shard_events(counts, world, edges_per_node)[rank]
    return slice_batch(batch, lo, hi)


def allreduce_sum_(flat_grad: torch.Tensor, group=None) -> float:
    """SUM all-reduce of the flat gradient buffer in place; returns the scale (1/world) the optimiser
    kernel must apply.  A no-op returning 1.0 outside a process group."""
    if group is None and not (dist.is_available() and dist.is_initialized()):
        return 1.0
    world = dist.get_world_size(group)
    if world == 1:
        return 1.0
    dist.all_reduce(flat_grad, op=dist.ReduceOp.SUM, group=group)
    return 1.0 / world
