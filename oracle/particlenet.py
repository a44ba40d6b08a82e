"""CPU restatement of the reference's ParticleNeT in eval mode (TEST INFRASTRUCTURE ONLY - see
oracle/__init__.py).

Follows /root/reference/src/graphnet/models/gnn/particlenet.py:176-256: DynEdgeConv blocks with
Linear - BatchNorm1d - activation MLPs and aggr="mean" (components/layers.py:62-76: EdgeConv, then
kNN on `x[:, features_subset]`), optional re-use of the input graph (`dynamic=False`), global
pooling (:224-238) and the readout Linear - activation - Dropout stack (:212-222; Dropout is the
identity in eval mode).  Pinned against the reference's own file run under oracle/refstub.py
(tests/golden/make_golden.py, case oscnext5_particlenet).
"""

from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import torch
from torch import Tensor

from . import ops

_POOL = {
    "min": lambda x, b, M: ops.scatter_min(x, b, M)[0],
    "max": lambda x, b, M: ops.scatter_max(x, b, M)[0],
    "sum": ops.scatter_sum,
    "mean": ops.scatter_mean,
}
DEFAULT_SIZES = [(64, 64, 64), (128, 128, 128), (256, 256, 256)]


def init_particlenet_state_dict(nb_inputs: int, sizes: Sequence[Tuple[int, ...]] = DEFAULT_SIZES,
                                readout: Sequence[int] = (256,), pools: Sequence[str] = ("mean",),
                                batchnorm: bool = True, seed: int = 0) -> Dict[str, Tensor]:
    """Parameters as torch creates them for the reference module under `torch.manual_seed(seed)`
    (construction order of particlenet.py:176-222), then non-trivial BatchNorm statistics so
    that the normalisation is exercised (drawn from the same generator, after all layers)."""
    torch.manual_seed(seed)
    sd: Dict[str, Tensor] = {}
    bns: List[str] = []
    lat = nb_inputs
    step = 3 if batchnorm else 2
    for i, s in enumerate(sizes):
        dims = [lat] + list(s)
        for j, (a, b) in enumerate(zip(dims[:-1], dims[1:])):
            lin = torch.nn.Linear(a * 2 if j == 0 else a, b)
            sd[f"_conv_layers.{i}.nn.{j * step}.weight"] = lin.weight.detach().clone()
            sd[f"_conv_layers.{i}.nn.{j * step}.bias"] = lin.bias.detach().clone()
            if batchnorm:
                bns.append(f"_conv_layers.{i}.nn.{j * step + 1}")
                sd[bns[-1] + ".weight"] = torch.ones(b)
                sd[bns[-1] + ".bias"] = torch.zeros(b)
                sd[bns[-1] + ".running_mean"] = torch.zeros(b)
                sd[bns[-1] + ".running_var"] = torch.ones(b)
                sd[bns[-1] + ".num_batches_tracked"] = torch.tensor(0)
        lat = dims[-1]
    dims = [lat * max(1, len(pools))] + list(readout)
    for j, (a, b) in enumerate(zip(dims[:-1], dims[1:])):
        lin = torch.nn.Linear(a, b)
        sd[f"_readout.{3 * j}.weight"] = lin.weight.detach().clone()
        sd[f"_readout.{3 * j}.bias"] = lin.bias.detach().clone()
    for name in bns:
        n = sd[name + ".weight"].shape[0]
        sd[name + ".weight"] = 0.5 + torch.rand(n)
        sd[name + ".bias"] = 0.2 * torch.randn(n)
        sd[name + ".running_mean"] = 0.3 * torch.randn(n)
        sd[name + ".running_var"] = 0.5 + torch.rand(n)
    return sd


def particlenet_forward(sd: Dict[str, Tensor], x: Tensor, edge_index: Tensor, batch: Tensor,
                        sizes: Sequence[Tuple[int, ...]] = DEFAULT_SIZES,
                        readout: Sequence[int] = (256,), pools: Optional[Sequence[str]] = ("mean",),
                        nb_neighbours: int = 16, features_subset=slice(0, 3), dynamic: bool = True,
                        batchnorm: bool = True, activation: str = "relu",
                        skip_readout: bool = False, return_intermediates: bool = False):
    F = torch.nn.functional
    act = torch.relu if activation == "relu" else F.gelu
    step = 3 if batchnorm else 2
    convs, graphs = [], [edge_index]
    ei = edge_index
    for i, s in enumerate(sizes):
        def nn_(t: Tensor, i=i, n=len(s)) -> Tensor:
            for j in range(n):
                p = f"_conv_layers.{i}.nn.{j * step}"
                t = F.linear(t, sd[p + ".weight"], sd[p + ".bias"])
                if batchnorm:
                    q = f"_conv_layers.{i}.nn.{j * step + 1}"
                    t = F.batch_norm(t, sd[q + ".running_mean"], sd[q + ".running_var"],
                                     sd[q + ".weight"], sd[q + ".bias"], False, 0.1, 1e-5)
                t = act(t)
            return t
        x = ops.edge_conv(x, ei, nn_, aggr="mean")
        new_ei = ops.knn_graph(x[:, features_subset], nb_neighbours, batch)
        if dynamic:
            ei = new_ei
        convs.append(x)
        graphs.append(ei)
    if not skip_readout:
        if pools:
            M = int(batch.max()) + 1
            x = torch.cat([_POOL[p](x, batch, M) for p in pools], dim=1)
        for j in range(len(readout)):
            x = act(F.linear(x, sd[f"_readout.{3 * j}.weight"], sd[f"_readout.{3 * j}.bias"]))
    if return_intermediates:
        return x, {"convs": convs, "graphs": graphs}
    return x
