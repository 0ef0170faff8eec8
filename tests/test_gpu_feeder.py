"""Packed ragged batches + double-buffered host feed (row f3; replaces model/data_loader.py:59-61,107-110 + PyG
collate): the feeder must deliver exactly the host bytes, in order, and training through it must equal training on
device-resident batches bit for bit."""
import pytest
import torch

from deepmetv2_b200 import synth
from deepmetv2_b200.feeder import HostFeeder, pack_events
from deepmetv2_b200.model import net as gnet
from deepmetv2_b200.trainer import GraphMETTrainer

pytestmark = pytest.mark.gpu
DEV = "cuda"


def test_feeder_delivers_host_batches_in_order():
    host = [synth.make_batch(3, "dytt", seed=50 + k, sizes=[120 + 7 * k, 90, 150 - 5 * k]).pin_memory() for k in range(7)]
    seen = 0
    for k, b in enumerate(HostFeeder(host, DEV, depth=2)):
        assert torch.equal(b.x.cpu(), host[k].x) and torch.equal(b.y.cpu(), host[k].y)
        assert torch.equal(b.ptr.cpu(), host[k].ptr)
        assert b.max_event_nodes == int((host[k].ptr[1:] - host[k].ptr[:-1]).max())
        seen += 1
    assert seen == len(host)
    assert list(HostFeeder([], DEV)) == []


def test_training_through_the_feeder_is_bit_identical():
    host = [synth.make_batch(2, "dytt", seed=60 + k, sizes=[200, 170 + k]).pin_memory() for k in range(5)]
    torch.manual_seed(1)
    a = gnet.Net(8, 3).to(DEV)
    b = gnet.Net(8, 3).to(DEV); b.load_state_dict(a.state_dict())
    ta, tb = GraphMETTrainer(a), GraphMETTrainer(b)
    la = ta.train_epoch_from_host(host)
    lb = [float(tb.train_step(h.to(DEV)).item()) for h in host]
    assert la == lb
    assert torch.equal(ta.flat, tb.flat)
    ea = ta.eval_epoch_from_host(host[:2])
    eb = [float(tb.eval_step(h.to(DEV))[1].item()) for h in host[:2]]
    assert ea == eb
