"""Where the HOST time of a training step goes (cProfile over a few steps of bench.py's C3 step)."""
import cProfile
import os
import pstats
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import graphnet_b200 as g  # noqa: E402
from graphnet_b200.synthetic import make_pulse_batch  # noqa: E402

dev = torch.device("cuda:0")
pb = make_pulse_batch(1024, 60.0, seed=1234)
x, batch, npulse = pb.x.to(dev), pb.batch.to(dev), pb.n_pulses.to(dev)
ei = g.knn_graph(x[:, :3].contiguous(), 8, batch, num_events=1024)
target = (torch.rand(1024) * 4).to(dev)
torch.manual_seed(0)
backbone = g.DynEdge(7, global_pooling_schemes=["min", "max", "mean", "sum"], precision="bf16")
task = g.EnergyReconstruction(hidden_size=backbone.nb_outputs, target_labels="energy",
                              loss_function=g.LogCoshLoss())
model = g.StandardModel(tasks=[task], graph_definition=object(), gnn=backbone,
                        optimizer_kwargs={"lr": 1e-3, "eps": 1e-3}).to(dev)
model.train()
backbone.validate_inputs = False
opt = model.configure_optimizers()["optimizer"]
data = g.Data(x=x, edge_index=ei, batch=batch, n_pulses=npulse, energy=target)


def step():
    opt.zero_grad(set_to_none=True)
    loss = model.training_step(data, 0)
    loss.backward()
    opt.step()


for _ in range(5):
    step()
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(11)]
t0 = time.perf_counter()
for i in range(10):
    ev[i].record()
    step()
ev[10].record()
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"host enqueue {1e3 * (t1 - t0) / 10:.2f} ms/step, with final sync {1e3 * (t2 - t0) / 10:.2f} ms/step; "
      f"device {ev[0].elapsed_time(ev[10]) / 10:.2f} ms/step")
import subprocess
print(subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_throttle_reasons.active",
                      "--format=csv,noheader"], capture_output=True, text=True).stdout.strip())
print("cpu:", os.cpu_count(), open("/proc/loadavg").read().strip())
pr = cProfile.Profile()
pr.enable()
for _ in range(5):
    step()
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(16)
st.sort_stats("cumtime").print_stats(45)
