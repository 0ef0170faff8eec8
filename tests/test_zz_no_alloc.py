"""gm_radius_graph_static — the radius graph through the library's static device scratch (no allocation inside the
call) — against the existing two-call entry points gm_radius_graph_count / gm_radius_graph_fill, through the C-ABI.
Integer work: the bar is bit-exact (rowptr, col, stats).  Collected last on purpose (file name).
The entry point has no counterpart in the reference."""
import pytest
import torch

from deepmetv2_b200 import ops, synth
from deepmetv2_b200._lib import check, lib, ptr, stream

pytestmark = pytest.mark.gpu
DEV = "cuda"
MAX_NODES, MAX_EVENTS = 65536, 64            # GM_STATIC_MAX_NODES / GM_STATIC_MAX_EVENTS of include/graphmet.h


def _static(etaphi, ptr_i32, r, max_nn, loop, capacity):
    N, B = etaphi.shape[0], ptr_i32.numel() - 1
    max_event = int((ptr_i32[1:] - ptr_i32[:-1]).max().item()) if B > 0 else 0
    # outputs are the caller's; pre-filled so that stale contents cannot pass for results
    rowptr = torch.full((N + 1,), -7, dtype=torch.int32, device=DEV)
    col = torch.full((max(capacity, 1),), -7, dtype=torch.int32, device=DEV)
    stats = torch.full((4,), -7, dtype=torch.int32, device=DEV)
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats()["allocation.all.allocated"]     # cumulative count: frees by the GC do not move it
    rc = lib().gm_radius_graph_static(ptr(etaphi), etaphi.stride(0), 2, ptr(ptr_i32), N, B, float(r), max_nn, int(loop), 0,
                                      max_event, ptr(rowptr), ptr(col), capacity, ptr(stats), stream())
    check(rc, "gm_radius_graph_static")
    torch.cuda.synchronize()
    assert torch.cuda.memory_stats()["allocation.all.allocated"] == before   # nothing came from torch's allocator either
    return rowptr, col, stats


def _case(n_events, shape, seed, r, max_nn, loop):
    g = synth.make_batch(n_events, shape, seed=seed).to(DEV)
    _, _, etaphi = ops.prepare_inputs(g.x)
    return etaphi.contiguous(), g.ptr


def test_static_scratch_matches_two_call_protocol_on_two_calls_in_a_row():
    # two DIFFERENT problems back to back: the second call finds the first one's contents in the shared scratch
    cases = [(_case(32, "dytt", 11, 0.4, 255, True), 0.4, 255, True),
             (_case(6, "highpu", 12, 0.4, 32, False), 0.4, 32, False)]
    for _ in range(2):                                      # and the same pair again
        for (etaphi, p), r, max_nn, loop in cases:
            ref = ops.build_graph(etaphi, r, p, loop=loop, max_num_neighbors=max_nn)    # _count + _fill, own workspace
            assert etaphi.shape[0] <= MAX_NODES and ref.num_edges > 0
            rowptr, col, stats = _static(etaphi, p, r, max_nn, loop, ref.num_edges)
            assert torch.equal(rowptr, ref.rowptr)
            assert torch.equal(col[: ref.num_edges], ref.col[: ref.num_edges])
            assert stats.tolist() == ref._stats.tolist()
            assert stats[0].item() == ref.num_edges == rowptr[-1].item()


def test_static_scratch_capped_rows_and_empty_input():
    (etaphi, p) = _case(4, "highpu", 13, 0.4, 16, True)    # the cap binds on most targets
    ref = ops.build_graph(etaphi, 0.4, p, loop=True, max_num_neighbors=16)
    assert ref.num_capped > 0
    rowptr, col, stats = _static(etaphi, p, 0.4, 16, True, ref.num_edges)
    assert torch.equal(rowptr, ref.rowptr) and torch.equal(col[: ref.num_edges], ref.col[: ref.num_edges])
    assert stats.tolist() == ref._stats.tolist()
    empty = torch.zeros(0, 2, device=DEV)
    p0 = torch.zeros(2, dtype=torch.int32, device=DEV)
    rowptr, _, stats = _static(empty, p0, 0.4, 16, True, 0)
    assert rowptr.tolist() == [0] and stats[0].item() == 0


def test_static_scratch_rejects_sizes_beyond_its_maximum(capfd):
    l = lib()
    one = torch.zeros(8, dtype=torch.int32, device=DEV)
    pos = torch.zeros(8, 2, device=DEV)
    for N, B in ((MAX_NODES + 1, 1), (16, MAX_EVENTS + 1)):
        rc = l.gm_radius_graph_static(ptr(pos), 2, 2, ptr(one), N, B, 0.4, 32, 0, 0, 16, ptr(one), ptr(one), 8, None,
                                      stream())
        assert rc == -2                                  # GM_ERR_CAPACITY
    assert "beyond the static scratch" in capfd.readouterr().err
    torch.cuda.synchronize()
