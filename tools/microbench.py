"""Per-call device timings of individual C-ABI entry points at a bench workload's size (GPU box).
    python tools/microbench.py [c2|c3|c5]
CUDA events on the launching stream, 5 warm-up + 20 timed calls, 256 MiB memset between calls (L2 flush)."""
import ctypes as C, os, sys
import torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from deepmetv2_b200 import ops, synth, _lib
from deepmetv2_b200._lib import lib, ptr, stream, check

def timeit(fn, flush, reps=20, warm=5):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2], ts[0]

def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
    n_ev, shape = {"c2": (32, "dytt"), "c3": (256, "highpu"), "c5": (32, "highpu")}[wl]
    b = synth.make_batch(n_ev, shape, seed=1000 * int(wl[1])).to("cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    l = lib()
    x = b.x
    N = x.shape[0]
    _, _, etaphi = ops.prepare_inputs(x)
    g = ops.build_graph(etaphi, 0.4, ptr_i32=b.ptr, loop=True, max_num_neighbors=255)
    E = g.num_edges
    print(f"{wl}: N={N} E={E} E/N={E / N:.1f}")
    men = int((b.ptr[1:] - b.ptr[:-1]).max())
    t = timeit(lambda: ops.build_graph(etaphi, 0.4, ptr_i32=b.ptr, loop=True, max_num_neighbors=255,
                                       max_event_nodes=men, edge_capacity=E), flush)
    print(f"radius graph (count+fill, no sync)      median {t[0]:8.1f} us  min {t[1]:8.1f} us")
    emb = torch.randn(N, 32, device="cuda")
    W = torch.randn(32, 64, device="cuda") * 0.1
    bias = torch.randn(32, device="cuda")
    out = torch.empty(N, 32, device="cuda")
    arg = torch.empty(N, 32, dtype=torch.int32, device="cuda")
    ws = torch.empty(l.gm_edgeconv_ws_bytes(N), dtype=torch.uint8, device="cuda")
    for name, prec, a in (("fp32 factored (node_prep + edge_aggr)", 0, None), ("fp32 factored + arg", 0, arg),
                          ("fp32 direct (per-edge FFMA)", 4, None), ("tf32 tcgen05", 1, None), ("bf16 tcgen05", 2, None)):
        fn = lambda: check(l.gm_edgeconv_forward(ptr(emb), ptr(g.rowptr), ptr(g.col), N, ptr(W), ptr(bias), 0, prec,
                                                 ptr(ws), ws.numel(), ptr(out), ptr(a), stream()))
        t = timeit(fn, flush)
        print(f"edgeconv {name:38s} median {t[0]:8.1f} us  min {t[1]:8.1f} us   {E / t[0] / 1e3:7.2f} Gedge/s")

if __name__ == "__main__":
    main()
