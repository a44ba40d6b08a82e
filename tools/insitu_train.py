"""In-step duration of every kernel of the C2 training step (torch.profiler / CUPTI), aggregated by
kernel name, plus the wall time of the step: how far the GPU time is from the step time."""
import collections
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import graphnet_b200 as g  # noqa: E402
from graphnet_b200.synthetic import make_pulse_batch  # noqa: E402

dev = torch.device("cuda:0")
pb = make_pulse_batch(1024, 60.0, seed=1234)
x, batch, npulse = pb.x.to(dev), pb.batch.to(dev), pb.n_pulses.to(dev)
ei = g.knn_graph(x[:, :3].contiguous(), 8, batch, num_events=1024)
target = torch.rand(1024, generator=torch.Generator().manual_seed(7)) * 4
torch.manual_seed(0)
backbone = g.DynEdge(7, global_pooling_schemes=["min", "max", "mean", "sum"], precision="bf16")
task = g.EnergyReconstruction(hidden_size=backbone.nb_outputs, target_labels="energy",
                              loss_function=g.LogCoshLoss())
model = g.StandardModel(tasks=[task], graph_definition=object(), gnn=backbone,
                        optimizer_kwargs={"lr": 1e-3, "eps": 1e-3}).to(dev)
model.train()
backbone.validate_inputs = False
opt = model.configure_optimizers()["optimizer"]
data = g.Data(x=x, edge_index=ei, batch=batch, n_pulses=npulse, energy=target.to(dev))


def step():
    opt.zero_grad(set_to_none=True)
    loss = model.training_step(data, 0)
    loss.backward()
    opt.step()
    return loss


for _ in range(5):
    step()
torch.cuda.synchronize()
STEPS = 5
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(STEPS):
        step()
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
agg = collections.defaultdict(lambda: [0, 0.0])
for e in evs:
    a = agg[e.name[:100]]
    a[0] += 1
    a[1] += e.time_range.elapsed_us()
tot = sum(v[1] for v in agg.values()) / STEPS
print(f"{len(evs) / STEPS:.0f} device activities per step; sum of kernels {tot:.0f} us/step; "
      f"wall {(evs[-1].time_range.end - evs[0].time_range.start) / STEPS:.0f} us/step")
for name, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:40]:
    print(f"{n / STEPS:6.1f} x {us / n:8.1f} us = {us / STEPS:8.1f} us/step  {name}")
