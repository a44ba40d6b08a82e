import torch, sys
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import oracle, graphnet_b200 as gb, graphnet_b200.ops as ops
from graphnet_b200.dynedge_bf16 import forward_bf16
from conftest import load_golden
DEV='cuda:0'
def rel(a,b): return float((a.double()-b.double()).norm()/b.double().norm())
import itertools
for name, prec in itertools.product(["upgrade5","prometheus50","oscnext5"], ["bf16","fp16"]):
    g = load_golden(name)
    spec = oracle.DynEdgeSpec(nb_inputs=g["x"].shape[1], **g["dyn_kwargs"])
    sd = oracle.init_dynedge_state_dict(spec, seed=int(g["seed"]))
    m = gb.DynEdge(g["x"].shape[1], precision=prec, **g["dyn_kwargs"]); m.load_state_dict(sd); m=m.to(DEV)
    data = gb.Data(x=g["x"].to(DEV), edge_index=g["edge_index"].to(DEV), batch=g["batch"].to(DEV), n_pulses=g["n_pulses"].to(DEV))
    out, inter = forward_bf16(m, data, return_intermediates=True)
    offs = inter["skip_offsets"]; skip = inter["skip"].float().cpu()
    print(name, prec)
    for i in range(4):
        c = skip[:, offs[i]:offs[i]+256]
        ei = ops.ell_to_coo(inter["graphs"][i]).cpu(); want = g[f"graph{i+1}"]
        same = (ei.shape == want.shape) and float((ei[0]==want[0]).float().mean())
        print(f"  conv{i} rel {rel(c, g[f'conv{i}']):.4f}  coords rel {rel(inter['coords'][i].cpu(), g[f'conv{i}'][:, :3]):.5f}  graph{i+1} same-edge frac {same}")
    print("  post", rel(inter["post"].cpu(), g["post"]), "pooled", rel(inter["pooled"].cpu(), g["pooled"]) if "pooled" in g else None, "out", rel(out.cpu(), g["out"]))
