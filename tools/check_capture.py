"""Does ops.run_captured replay the fp32 forward as a graph? (prints the driver's answer)"""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import graphnet_b200 as g
from graphnet_b200 import _lib, ops
from graphnet_b200.synthetic import make_pulse_batch
lib = _lib.load()
orig = lib.gn_graph_capture_launch
def wrapped(key, st):
    rc = orig(key, st)
    print("capture_launch rc", rc, lib.gn_last_error().decode() if rc else "")
    return rc
lib.gn_graph_capture_launch = wrapped
dev = torch.device("cuda:0")
pb = make_pulse_batch(256, 60.0, seed=1)
m = g.DynEdge(7, global_pooling_schemes=["min", "max", "mean", "sum"], precision="fp32").to(dev).eval()
m.validate_inputs = False
d = g.Data(x=pb.x.to(dev), batch=pb.batch.to(dev), n_pulses=pb.n_pulses.to(dev))
for i in range(3):
    out = m(d)
torch.cuda.synchronize()
import time
for tag, env in (("graph", None), ("direct", "1")):
    if env: os.environ["GN_NO_GRAPH"] = env
    lib.gn_graph_capture_launch = orig
    for _ in range(5): m(d)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(50): m(d)
    torch.cuda.synchronize(); print(tag, (time.perf_counter() - t0) / 50 * 1e3, "ms/forward")
