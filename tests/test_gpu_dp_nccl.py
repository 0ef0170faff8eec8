"""Two NCCL ranks through ``GraphMETTrainer(process_group=...)`` -- the code path SCALE times (DP had
only been exercised with the oracle on gloo).  Needs 2 GPUs (``gpurun --gpus 2``); skipped on a 1-GPU box.

Checked, for the fused peer-memory exchange (csrc/comm.cu) and for the torch.distributed fallback:
  * after every step all ranks hold BIT-IDENTICAL parameters and optimizer moments;
  * the update is the gradient of the MEAN loss over all events of the step even though the ranks hold different
    numbers of events (3 + 2): exp_avg after step 1 == (1 - beta1) * sum_r B_r g_r / sum_r B_r with g_r from
    single-process trainers on the shards;
  * a rank with an EMPTY shard (B < world) joins the exchange with a zero gradient instead of dead-locking;
  * the returned loss is the mean over all ranks' events.
"""
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, mode, out_dir):
    import torch.distributed as dist
    from deepmetv2_b200 import dp, synth
    from deepmetv2_b200.model import net as gnet
    from deepmetv2_b200.synth import EventBatch
    from deepmetv2_b200.trainer import GraphMETTrainer
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        torch.manual_seed(0)
        model = gnet.Net(8, 3).to(dev)
        fused = mode != "nccl"
        tr = GraphMETTrainer(model, lr=1e-3, process_group=dist.group.WORLD, fused_allreduce=fused,
                             overlap_exchange=(mode == "overlap"))
        assert (tr.comm is not None) == fused, "CUDA IPC peer mapping failed on this box"
        full = synth.make_batch(5, "dytt", seed=90, sizes=[230, 180, 260, 150, 205])
        shards = [(0, 3), (3, 5)]                               # 3 + 2 events: different B per rank
        record = {}
        for step in range(3):
            if step < 2:
                lo, hi = shards[rank]
                b = dp.slice_batch(full, lo, hi)
            else:                                               # step 3: rank 1 has NO events
                b = dp.slice_batch(full, 0, 5) if rank == 0 else dp.slice_batch(full, 5, 5)
            g = EventBatch(b.x.to(dev), b.y.to(dev), None, b.ptr.to(dev).int())
            loss = tr.train_step(g, max_event_nodes=300)
            if mode == "overlap":
                assert tr._pending is not None          # the update of this step waits for the next step / flush()
            tr.flush()
            torch.cuda.synchronize()
            record[f"loss{step}"] = loss.item()
            record[f"flat{step}"] = tr.flat.cpu().clone()
            record[f"m{step}"] = tr.exp_avg.cpu().clone()
            record[f"v{step}"] = tr.exp_avg_sq.cpu().clone()
        tr.sync_bn_buffers()
        record["bn"] = {k: v.cpu().clone() for k, v in model.state_dict().items() if "running" in k}
        torch.save(record, os.path.join(out_dir, f"rank{rank}.pt"))
        dist.barrier()
        tr.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["overlap", "fused", "nccl"])
def test_two_rank_training_is_replica_exact_and_globally_weighted(tmp_path, mode):
    """overlap: gm_exchange_post at the end of a step, gm_exchange_apply_adamw behind the next step's graph build (here:
    in flush()); fused: both in one kernel; nccl: torch.distributed.all_reduce + gm_adamw_step_ex."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    from deepmetv2_b200 import dp, synth
    from deepmetv2_b200.model import net as gnet
    from deepmetv2_b200.synth import EventBatch
    from deepmetv2_b200.trainer import GraphMETTrainer
    mp.spawn(_worker, args=(2, _free_port(), mode, str(tmp_path)), nprocs=2, join=True)
    r0 = torch.load(tmp_path / "rank0.pt", weights_only=False)
    r1 = torch.load(tmp_path / "rank1.pt", weights_only=False)
    for step in range(3):
        for key in ("flat", "m", "v"):
            assert torch.equal(r0[f"{key}{step}"], r1[f"{key}{step}"]), (key, step)
        assert r0[f"loss{step}"] == r1[f"loss{step}"]
    for k in r0["bn"]:
        assert torch.equal(r0["bn"][k], r1["bn"][k]), k          # after sync_bn_buffers: rank 0's statistics everywhere
    # expected first update from single-process trainers on the two shards (same init)
    full = synth.make_batch(5, "dytt", seed=90, sizes=[230, 180, 260, 150, 205])
    grads, losses, counts = [], [], []
    for lo, hi in [(0, 3), (3, 5)]:
        torch.manual_seed(0)
        m = gnet.Net(8, 3).to("cuda")
        t = GraphMETTrainer(m, lr=1e-3)
        b = dp.slice_batch(full, lo, hi)
        l = t.train_step(EventBatch(b.x.cuda(), b.y.cuda(), None, b.ptr.cuda().int()), max_event_nodes=300)
        grads.append(t.flat_grad.cpu().double().clone()); losses.append(l.item()); counts.append(hi - lo)
    want = (counts[0] * grads[0] + counts[1] * grads[1]) / sum(counts)
    got = r0["m0"].double() / (1.0 - 0.9)
    assert (got - want).abs().max().item() <= 1e-5 * want.abs().max().item()
    want_loss = (counts[0] * losses[0] + counts[1] * losses[1]) / sum(counts)
    assert abs(r0["loss0"] - want_loss) <= 1e-6 * abs(want_loss)


def _worker_overlap_steps(rank, world, port, overlap, out_dir):
    """Several back-to-back steps WITHOUT flushing in between (the deferred update runs inside the next step)."""
    import torch.distributed as dist
    from deepmetv2_b200 import synth
    from deepmetv2_b200.model import net as gnet
    from deepmetv2_b200.trainer import GraphMETTrainer
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        torch.manual_seed(0)
        model = gnet.Net(8, 3).to(dev)
        tr = GraphMETTrainer(model, lr=1e-3, process_group=dist.group.WORLD, overlap_exchange=overlap)
        batches = [synth.make_batch(3, "dytt", seed=200 + 10 * rank + i, sizes=[220, 260, 180]).to(dev) for i in range(3)]
        for step in range(6):
            tr.train_step(batches[step % 3], max_event_nodes=300)
        tr.flush()
        torch.cuda.synchronize()
        torch.save({"flat": tr.flat.cpu().clone(), "loss": tr.global_loss.item()}, os.path.join(out_dir, f"o{int(overlap)}_rank{rank}.pt"))
        dist.barrier()
        tr.close()
    finally:
        dist.destroy_process_group()


def test_deferred_exchange_equals_fused_exchange_after_six_steps(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    for overlap in (True, False):
        mp.spawn(_worker_overlap_steps, args=(2, _free_port(), overlap, str(tmp_path)), nprocs=2, join=True)
    a0, a1 = (torch.load(tmp_path / f"o1_rank{r}.pt", weights_only=False) for r in (0, 1))
    b0 = torch.load(tmp_path / "o0_rank0.pt", weights_only=False)
    assert torch.equal(a0["flat"], a1["flat"])                    # replicas identical
    assert torch.equal(a0["flat"], b0["flat"])                    # and bit-identical to the one-kernel exchange
    assert a0["loss"] == b0["loss"]
