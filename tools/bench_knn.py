"""Time the k-NN graph search (csrc/knn.cu) on the bench workloads' shapes; CUDA events, L2 flushed between launches."""
import json, sys, time
import torch
from deepmetv2_b200 import ops, synth
from deepmetv2_b200._lib import lib, ptr, stream, check

def run(workload, B, shape, dim, k, loop, steps=10):
    b = synth.make_batch(B, shape, seed=2000)
    N = b.x.shape[0]
    torch.manual_seed(0)
    x = torch.randn(N, dim, device="cuda")
    p = b.ptr.to("cuda").int()
    l = lib()
    ws = torch.empty(l.gm_knn_graph_ws_bytes(N, k, int(loop)), dtype=torch.uint8, device="cuda")
    rowptr = torch.empty(N + 1, dtype=torch.int32, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ts = []
    for it in range(steps + 3):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        check(l.gm_knn_graph_count(ptr(x), x.stride(0), dim, ptr(p), N, B, k, int(loop), ptr(ws), ws.numel(), ptr(rowptr),
                                   stream()), "knn")
        e1.record(); torch.cuda.synchronize()
        if it >= 3: ts.append(e0.elapsed_time(e1))
    ms = sum(ts) / len(ts)
    sizes = (b.ptr[1:] - b.ptr[:-1]).double()
    pairs = float((sizes * sizes).sum())
    print(json.dumps({"workload": workload, "N": N, "B": B, "dim": dim, "k": k, "loop": loop, "ms": ms,
                      "events_per_s": B / ms * 1e3, "pair_dims_per_s": pairs * dim / ms * 1e3,
                      "fp32_flops": 3 * pairs * dim / ms * 1e3}))

if __name__ == "__main__":
    run("c2-shape knn(emb32, k=20, loop)", 32, "dytt", 32, 20, True)
    run("c2-shape knn(latent64, k=8, !loop)", 32, "dytt", 64, 8, False)
    run("c5-shape knn(emb32, k=20, loop)", 32, "highpu", 32, 20, True)
