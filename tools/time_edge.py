#!/usr/bin/env python
"""Time the fused EdgeConv kernel alone (CUDA events, flushed L2) on C2 / C4 shapes.

    [GN_B200_DEBUG_LIB=1 GN_EDGE_DEBUG=1] python tools/time_edge.py [c2|c4]

Prints one JSON line per conv layer shape; with GN_EDGE_DEBUG=1 also the kernel's wait-time
counters per step (cycles), read through the diagnostic export gn_debug_edge_counters of the profiling build
(`python __graft_entry__.py --debug`).
"""
import ctypes
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import graphnet_b200 as g  # noqa: E402
from graphnet_b200 import _lib  # noqa: E402
from graphnet_b200.dynedge_bf16 import time_edge_kernel  # noqa: E402
from graphnet_b200.synthetic import make_pulse_batch  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
dev = torch.device("cuda:0")
if wl == "c4":
    pb = make_pulse_batch(256, 500.0, detector="upgrade", seed=1234)
    k, pools = 16, ["min", "max", "mean"]
else:
    pb = make_pulse_batch(1024, 60.0, seed=1234)
    k, pools = 8, ["min", "max", "mean", "sum"]
x, batch, npl = pb.x.to(dev), pb.batch.to(dev), pb.n_pulses.to(dev)
model = g.DynEdge(x.shape[1], nb_neighbours=k, global_pooling_schemes=pools, precision="bf16").to(dev)
data = g.Data(x=x, batch=batch, n_pulses=npl)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
lib = _lib.load()
dbg = os.environ.get("GN_EDGE_DEBUG") is not None
for layer in (1, 0):
    if dbg:
        buf = (ctypes.c_ulonglong * 8)()
        lib.gn_debug_edge_counters(buf)   # clear
    reps = 10
    ms, flops, name = time_edge_kernel(model, data, flush, layer=layer, reps=reps)
    line = {"workload": wl, "layer": layer, "us": ms * 1e3, "executed_tflops": flops / (ms * 1e-3) / 1e12,
            "kernel": name.split("(")[0].strip()}
    if dbg:
        lib.gn_debug_edge_counters(buf)
        c = list(buf)
        steps = max(c[7], 1)
        line["per_step_clk"] = {"bld_wait_h_empty": c[2] / steps / 32, "mma_wait_h_full": c[3] / steps,
                                "mma_wait_d_empty": c[4] / steps, "epi_wait_d_full": c[5] / steps / 16, "epi_ld_phase": c[0] / steps / 16,
                                "total": c[6] / steps, "steps": c[7],
                                "loop_clk_per_launch": c[6] / (74 * (reps + 3)),
                                "loop_clk_max_cluster": c[1], "raw": c}
        tb = (ctypes.c_ulonglong * (31 * 256 + 4))()
        lib.gn_debug_edge_trace(tb, 31 * 256 + 4)
        line["setup_clk_max"] = tb[31 * 256]
        line["kernel_clk_max"] = tb[31 * 256 + 1]
    print(json.dumps(line), flush=True)
