#!/usr/bin/env python
"""Timeline of cluster 0 of the fused EdgeConv kernel (profiling build).

    GN_B200_DEBUG_LIB=1 GN_EDGE_DEBUG=1 python tools/trace_edge.py [c2|c4] [layer] [first_step]

Per step and warp role, cycles since the set-up barrier: MMA issuer (h tile ready, accumulator
free, issued), epilogue warps (accumulator full, released, stored), builders (task start, loads
issued, h tile free, arrived).
"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import graphnet_b200 as g  # noqa: E402
from graphnet_b200 import _lib  # noqa: E402
from graphnet_b200.dynedge_bf16 import time_edge_kernel  # noqa: E402
from graphnet_b200.synthetic import make_pulse_batch  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
layer = int(sys.argv[2]) if len(sys.argv) > 2 else 1
s0 = int(sys.argv[3]) if len(sys.argv) > 3 else 24
dev = torch.device("cuda:0")
pb = make_pulse_batch(256, 500.0, detector="upgrade", seed=1234) if wl == "c4" else \
    make_pulse_batch(1024, 60.0, seed=1234)
k, pools = (16, ["min", "max", "mean"]) if wl == "c4" else (8, ["min", "max", "mean", "sum"])
x, batch, npl = pb.x.to(dev), pb.batch.to(dev), pb.n_pulses.to(dev)
model = g.DynEdge(x.shape[1], nb_neighbours=k, global_pooling_schemes=pools, precision="bf16").to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
lib = _lib.load()
time_edge_kernel(model, g.Data(x=x, batch=batch, n_pulses=npl), flush, layer=layer, reps=2)
W = 2 * 32 * 64 * 4
buf = (ctypes.c_ulonglong * W)()
lib.gn_debug_edge_trace(buf, W)
tr = torch.tensor(list(buf), dtype=torch.int64).reshape(2, 32, 64, 4)


def row(rank, warp, step):
    return [int(v) for v in tr[rank, warp, step]]


for step in (s0, s0 + 1):   # per-warp detail of two steps
    for rank in (0, 1):
        for w in range(8):
            print(f"detail step {step} r{rank} epi warp {w}: {row(rank, w, step)}")
        for w in range(8, 24):
            print(f"detail step {step} r{rank} bld warp {w}: {row(rank, w, step)}")
for step in range(s0, s0 + 6):
    m = row(0, 25, step)
    print(f"step {step}: MMA wait-start {m[0]} h_full {m[1]} d_empty {m[2]} issued {m[3]}")
    for rank in (0, 1):
        e = [row(rank, w, step) for w in range(8)]
        print(f"   epi r{rank}: d_full-wait-start min/max {min(v[0] for v in e)}/{max(v[0] for v in e)}"
              f"  d_full {min(v[1] for v in e)}/{max(v[1] for v in e)}"
              f"  released {min(v[2] for v in e)}/{max(v[2] for v in e)}"
              f"  stored {min(v[3] for v in e)}/{max(v[3] for v in e)}")
        b = [row(rank, w, step) for w in range(8, 24) if row(rank, w, step)[3] > 0]
        if b:
            print(f"   bld r{rank}: {len(b)} tasks  start {min(v[0] for v in b)}/{max(v[0] for v in b)}"
                  f"  loads-issued {min(v[1] for v in b)}/{max(v[1] for v in b)}"
                  f"  h_empty {min(v[2] for v in b)}/{max(v[2] for v in b)}"
                  f"  arrived {min(v[3] for v in b)}/{max(v[3] for v in b)}")
