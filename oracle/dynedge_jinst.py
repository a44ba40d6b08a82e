"""CPU restatement of the reference's DynEdgeJINST (TEST INFRASTRUCTURE ONLY - see oracle/__init__.py).

Follows /root/reference/src/graphnet/models/gnn/dynedge_jinst.py:19-160: four DynEdgeConv layers
(Linear-LeakyReLU-Linear-LeakyReLU, aggr="add", k=8, kNN on latent columns 0-2 after every layer,
components/layers.py:62-76), skip concatenation, nn1-LeakyReLU-nn2, max/min/sum/mean pooling,
the xyzt homophily of the INPUT graph (models/utils.py:17-33) and the raw pulse count appended,
LeakyReLU-nn3-LeakyReLU.  Pinned against the reference's own file run under oracle/refstub.py
(tests/golden/make_golden.py, case oscnext5_jinst).
"""

from __future__ import annotations

from typing import Dict

import torch
from torch import Tensor

from . import ops


def init_jinst_state_dict(nb_inputs: int, layer_size_scale: int = 4, seed: int = 0) -> Dict[str, Tensor]:
    """Parameters as torch creates them for the reference module under `torch.manual_seed(seed)`
    (same construction order: conv_add1..4, nn1, nn2, nn3)."""
    c = layer_size_scale
    l1, l2, l3, l4, l5, l6 = nb_inputs, c * 32, c * 64, c * 84, c * 64, c * 32
    torch.manual_seed(seed)
    sd: Dict[str, Tensor] = {}

    def lin(name, i, o):
        m = torch.nn.Linear(i, o)
        sd[name + ".weight"] = m.weight.detach().clone()
        sd[name + ".bias"] = m.bias.detach().clone()

    lin("conv_add1.nn.0", 2 * l1, l2)
    lin("conv_add1.nn.2", l2, l3)
    for i in (2, 3, 4):
        lin(f"conv_add{i}.nn.0", 2 * l3, l4)
        lin(f"conv_add{i}.nn.2", l4, l3)
    lin("nn1", 4 * l3 + l1, l4)
    lin("nn2", l4, l5)
    lin("nn3", 4 * l5 + 5, l6)
    return sd


def jinst_forward(sd: Dict[str, Tensor], x: Tensor, edge_index: Tensor, batch: Tensor,
                  n_pulses: Tensor, return_intermediates: bool = False):
    lrelu = torch.nn.functional.leaky_relu
    F = torch.nn.functional
    B = int(n_pulses.reshape(-1).shape[0])
    h = [ops.homophily(edge_index, x[:, i], batch) for i in range(4)]   # x, y, z, t
    feats = [x]
    cur, ei = x, edge_index
    graphs = [edge_index]
    for i in (1, 2, 3, 4):
        w0, b0 = sd[f"conv_add{i}.nn.0.weight"], sd[f"conv_add{i}.nn.0.bias"]
        w2, b2 = sd[f"conv_add{i}.nn.2.weight"], sd[f"conv_add{i}.nn.2.bias"]
        cur = ops.edge_conv(cur, ei, lambda t: lrelu(F.linear(lrelu(F.linear(t, w0, b0)), w2, b2)),
                            "add")
        ei = ops.knn_graph(cur[:, 0:3], 8, batch)
        feats.append(cur)
        graphs.append(ei)
    z = torch.cat(feats, dim=1)
    z = F.linear(lrelu(F.linear(z, sd["nn1.weight"], sd["nn1.bias"])), sd["nn2.weight"], sd["nn2.bias"])
    a = ops.scatter_max(z, batch, dim_size=B)[0]
    b = ops.scatter_min(z, batch, dim_size=B)[0]
    c = ops.scatter_sum(z, batch, dim_size=B)
    d = ops.scatter_mean(z, batch, dim_size=B)
    pooled = torch.cat([a, b, c, d, h[3].reshape(-1, 1), h[0].reshape(-1, 1), h[1].reshape(-1, 1),
                        h[2].reshape(-1, 1), n_pulses.reshape(-1, 1)], dim=1)
    out = lrelu(F.linear(lrelu(pooled), sd["nn3.weight"], sd["nn3.bias"]))
    if return_intermediates:
        return out, {"convs": feats[1:], "graphs": graphs, "pooled": pooled, "node": z}
    return out
