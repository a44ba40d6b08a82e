"""Where the HOST time of an inference step goes (cProfile over bench.py's C2 step)."""
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
torch.manual_seed(0)
backbone = g.DynEdge(7, global_pooling_schemes=["min", "max", "mean", "sum"], precision="bf16")
task = g.EnergyReconstruction(hidden_size=backbone.nb_outputs, target_labels="energy",
                              loss_function=g.LogCoshLoss())
model = g.StandardModel(tasks=[task], graph_definition=object(), gnn=backbone).to(dev)
model.inference()
model.train(False)
backbone.validate_inputs = False
data = g.Data(x=x, edge_index=ei, batch=batch, n_pulses=npulse)


def step():
    return model(data)[0]


for _ in range(10):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(50):
    step()
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"host enqueue {1e3 * (t1 - t0) / 50:.3f} ms/step, with final sync {1e3 * (t2 - t0) / 50:.3f} ms/step")
# host only: how fast can the host enqueue when it never waits for the device? (queue depth is
# bounded, so measure a short burst after a sync)
torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(3):
    step()
t1 = time.perf_counter()
print(f"burst of 3 steps: {1e3 * (t1 - t0) / 3:.3f} ms/step host time")
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(20):
    step()
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(28)
