"""Time gn_knn_graph alone (CUDA events, median of reps) on the shapes of BASELINE configs 2, 4, 5.

    python tools/bench_knn.py            # one JSON line per shape

Algorithmic bytes: 4*3*N + 4*(B+1) in, 4*N*(k+2) out.  The kernel is
instruction bound, so pair-evaluations per second (sum n_e^2 / time) is printed as well.
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import graphnet_b200.ops as ops  # noqa: E402


def run(B, n, k, fixed, reps=20):
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(7)
    if fixed:
        counts = torch.full((B,), int(n))
    else:
        counts = torch.poisson(torch.full((B,), float(n)), generator=g).clamp(min=2).long()
    N = int(counts.sum())
    x = torch.randn(N, 3, generator=g)
    dup = torch.rand(N, generator=g) < 0.3          # ~30 % of pulses share a sensor position
    src = (torch.arange(N) - 1).clamp(min=0)
    x[dup] = x[src[dup]]
    ptr = torch.zeros(B + 1, dtype=torch.int32)
    ptr[1:] = counts.cumsum(0).int()
    x, ptr = x.to(dev), ptr.to(dev)
    for _ in range(3):
        ops.knn_graph_ell(x, k, ptr, [0, 1, 2])
    torch.cuda.synchronize()
    # ten calls replayed from one CUDA graph per timing: the kernel, not the Python caller
    inner = 10
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        for _ in range(inner):
            keep = ops.knn_graph_ell(x, k, ptr, [0, 1, 2])
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        graph.replay()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3 / inner)
    ts.sort()
    us = ts[len(ts) // 2]
    pairs = float((counts.double() ** 2).sum())
    nbytes = 12 * N + 4 * (B + 1) + 4 * N * (k + 2)
    print(json.dumps({"events": B, "pulses": n, "k": k, "nodes": N, "us": round(us, 1),
                      "gpairs_per_s": round(pairs / us / 1e3, 1),
                      "algorithmic_GBps": round(nbytes / us / 1e3, 1)}))


if __name__ == "__main__":
    run(1024, 60, 8, False)        # C2
    run(256, 500, 16, False)       # C4
    for n in (16, 64, 128, 256, 512, 1024, 2048):   # C5 sweep points, ~0.5 M nodes
        run(max(1, 524288 // n), n, 8, True)
