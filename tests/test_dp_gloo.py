"""N>1 host logic on CPU: event partitioner + gradient all-reduce over gloo (world_size 2).

The CUDA kernels cannot run here, so the per-rank forward/backward is the oracle's; what is under test
is deepmetv2_b200.dp (partition, slicing, collective, 1/world scaling) — the code the GPU trainer calls.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from deepmetv2_b200 import dp, synth


def test_shard_events_balanced_and_contiguous():
    rng = np.random.default_rng(0)
    counts = torch.from_numpy(rng.integers(800, 3000, 64))
    for world in (1, 2, 3, 4, 8):
        blocks = dp.shard_events(counts, world)
        assert blocks[0][0] == 0 and blocks[-1][1] == 64
        assert all(a[1] == b[0] for a, b in zip(blocks, blocks[1:]))
        loads = [int(counts[lo:hi].sum()) for lo, hi in blocks]
        assert all(hi > lo for lo, hi in blocks)
        assert max(loads) - min(loads) <= 3000, loads        # within one event of perfect balance


def test_shard_events_edge_cases():
    assert dp.shard_events([5], 1) == [(0, 1)]
    assert dp.shard_events([], 2) == [(0, 0), (0, 0)]
    b = dp.shard_events([7, 7], 4)                            # fewer events than ranks: empty tails
    assert b[0][0] == 0 and b[-1][1] == 2 and sum(hi - lo for lo, hi in b) == 2
    b = dp.shard_events([1, 1, 1, 1000], 2)                   # a heavy last event still leaves rank 1 non-empty
    assert b == [(0, 3), (3, 4)]
    q = dp.shard_events([100] * 8, 2, edges_per_node=25.0)
    assert q == [(0, 4), (4, 8)]


def test_slice_batch_rebases():
    b = synth.make_batch(5, "tiny", seed=3)
    s = dp.slice_batch(b, 2, 5)
    assert s.ptr[0] == 0 and s.num_graphs == 3
    assert int(s.ptr[-1]) == s.x.shape[0] == int(b.ptr[5] - b.ptr[2])
    assert int(s.batch.min()) == 0 and int(s.batch.max()) == 2
    assert torch.equal(s.x, b.x[int(b.ptr[2]):int(b.ptr[5])])
    assert torch.equal(s.y, b.y[2:5])


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_grad(model, shard):
    from oracle import graphmet_oracle as O
    xc, xk, etaphi = O.prepare_inputs(shard.x)
    ei = O.radius_graph(etaphi, 0.4, shard.batch, loop=True, max_num_neighbors=255)
    model.zero_grad()
    loss = O.loss_fn(model(xc, xk, ei, shard.batch), shard.x, shard.y, shard.batch)
    loss.backward()
    return torch.cat([p.grad.reshape(-1) for p in model.parameters()]), loss.detach()


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import graphmet_oracle as O
        torch.manual_seed(0)
        torch.set_num_threads(1)
        model = O.Net(8, 3)
        model.train()                                          # per-rank BatchNorm statistics (DDP semantics)
        full = synth.make_batch(6, "tiny", seed=11)
        shard = dp.shard_batch(full, rank, world)
        flat, loss = _rank_grad(model, shard)
        scale = dp.allreduce_sum_(flat)
        torch.save({"grad": flat * scale, "loss": loss, "events": shard.num_graphs},
                   os.path.join(out_dir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_gradient_allreduce_matches_manual_average(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    got = [torch.load(tmp_path / f"rank{r}.pt") for r in range(world)]
    assert torch.equal(got[0]["grad"], got[1]["grad"])        # every rank holds the same averaged gradient
    assert got[0]["events"] + got[1]["events"] == 6

    from oracle import graphmet_oracle as O
    torch.manual_seed(0)
    torch.set_num_threads(1)
    full = synth.make_batch(6, "tiny", seed=11)
    grads = []
    for r in range(world):
        model = O.Net(8, 3)                                    # same seed => same init on every "rank"
        torch.manual_seed(0)
        model = O.Net(8, 3)
        model.train()
        grads.append(_rank_grad(model, dp.shard_batch(full, r, world))[0])
    want = (grads[0] + grads[1]) * 0.5
    assert torch.allclose(got[0]["grad"], want, rtol=0, atol=0)
    assert dp.allreduce_sum_(torch.zeros(3)) == 1.0            # outside a group: no-op


def test_pack_events_and_shard_round_trip(tmp_path):
    """Row f3 host logic: per-event tensors -> packed batch (x, y, ptr, batch) -> shard file -> the same batch."""
    import torch
    from deepmetv2_b200.feeder import pack_events, save_shard, load_shard
    torch.manual_seed(0)
    xs = [torch.randn(n, 11) for n in (5, 0, 3, 7)]
    ys = [torch.randn(11) for _ in xs]
    b = pack_events(xs, ys, pin=False)
    assert b.ptr.tolist() == [0, 5, 5, 8, 15] and b.x.shape == (15, 11) and b.y.shape == (4, 11)
    assert b.batch.tolist() == [0] * 5 + [2] * 3 + [3] * 7
    assert torch.equal(b.x[5:8], xs[2]) and torch.equal(b.y[3], ys[3])
    path = str(tmp_path / "shard.npz")
    save_shard(path, b)
    c = load_shard(path, pin=False)
    assert torch.equal(c.x, b.x) and torch.equal(c.y, b.y) and torch.equal(c.ptr, b.ptr) and torch.equal(c.batch, b.batch)
