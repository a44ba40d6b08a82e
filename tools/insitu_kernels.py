"""In-step duration of every kernel of the C2 inference forward (torch.profiler / CUPTI): unlike the
ncu launch list (cold caches, serialised) these are the times the kernels take inside the running
step, with the caches in the state the previous kernel left them.

    python tools/insitu_kernels.py [events] [mean_pulses] [precision] [c2|c4]
"""
import collections
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import graphnet_b200 as g  # noqa: E402
from graphnet_b200.synthetic import make_pulse_batch  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
mean = float(sys.argv[2]) if len(sys.argv) > 2 else 60.0
prec = sys.argv[3] if len(sys.argv) > 3 else "bf16"
c4 = len(sys.argv) > 4 and sys.argv[4] == "c4"
k = 16 if c4 else 8
dev = torch.device("cuda:0")
pb = make_pulse_batch(B, mean, seed=1234, detector="upgrade" if c4 else "icecube86")
x, batch, npulse = pb.x.to(dev), pb.batch.to(dev), pb.n_pulses.to(dev)
ei = g.knn_graph(x[:, :3].contiguous(), k, batch, num_events=B)
torch.manual_seed(0)
pools = ["min", "max", "mean"] if c4 else ["min", "max", "mean", "sum"]
m = g.DynEdge(x.shape[1], nb_neighbours=k, global_pooling_schemes=pools, precision=prec).to(dev).eval()
m.validate_inputs = False
data = g.Data(x=x, edge_index=ei, batch=batch, n_pulses=npulse)
for _ in range(10):
    m(data)
torch.cuda.synchronize()
STEPS = 20
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(STEPS):
        m(data)
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
per_step = len(evs) // STEPS
seq = collections.OrderedDict()
for i, e in enumerate(evs[: per_step * STEPS]):
    k = (i % per_step, e.name[:90])
    seq.setdefault(k, []).append(e.time_range.elapsed_us())
tot = 0.0
print(f"{per_step} device activities per step, N={x.shape[0]} pulses")
for (i, name), v in seq.items():
    avg = sum(v) / len(v)
    tot += avg
    print(f"{i:3d} {avg:8.1f} us  {name}")
t0 = evs[0].time_range.start
t1 = evs[per_step * STEPS - 1].time_range.end
print(f"sum of kernels {tot:.1f} us/step; wall {(t1 - t0) / STEPS:.1f} us/step")
