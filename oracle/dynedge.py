"""Oracle restatement of `graphnet.models.gnn.DynEdge` (CPU, torch).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Follows /root/reference/src/graphnet/models/gnn/dynedge.py:
  ctor defaults and checks      :25-182
  `_construct_layers`           :184-250   (parameter names / shapes)
  `_global_pooling`             :252-265
  `_calculate_global_variables` :267-294
  `forward`                     :296-350
and models/components/layers.py:62-76 (`DynEdgeConv.forward`: EdgeConv first,
then kNN on `x[:, features_subset]`), models/utils.py:17-33 (xyzt homophily).

Parameters are addressed by the reference's `state_dict` keys
(`_conv_layers.{i}.nn.{j}.weight`, `_post_processing.{j}.weight`,
`_readout.{j}.weight`, SURVEY.md section 5 checkpoint row).
"""

from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple, Union

import torch
from torch import Tensor

from . import ops

_POOL = {
    "min": lambda x, b, M: ops.scatter_min(x, b, M)[0],
    "max": lambda x, b, M: ops.scatter_max(x, b, M)[0],
    "sum": ops.scatter_sum,
    "mean": ops.scatter_mean,
}


@dataclass
class DynEdgeSpec:
    """Constructor arguments of the reference's DynEdge (dynedge.py:25-39)."""

    nb_inputs: int
    nb_neighbours: int = 8
    features_subset: Union[List[int], slice, None] = None
    dynedge_layer_sizes: Optional[List[Tuple[int, ...]]] = None
    post_processing_layer_sizes: Optional[List[int]] = None
    readout_layer_sizes: Optional[List[int]] = None
    global_pooling_schemes: Union[str, List[str], None] = None
    add_global_variables_after_pooling: bool = False
    activation_layer: Optional[str] = None
    add_norm_layer: bool = False
    skip_readout: bool = False
    # derived
    nb_global_variables: int = field(init=False)

    def __post_init__(self) -> None:
        if self.features_subset is None:
            self.features_subset = slice(0, 3)  # dynedge.py:79-80
        if self.dynedge_layer_sizes is None:  # dynedge.py:83-101
            self.dynedge_layer_sizes = [(128, 256), (336, 256), (336, 256), (336, 256)]
        if self.post_processing_layer_sizes is None:  # :113-118
            self.post_processing_layer_sizes = [336, 256]
        if self.readout_layer_sizes is None:  # :126-130
            self.readout_layer_sizes = [128]
        if isinstance(self.global_pooling_schemes, str):  # :138-139
            self.global_pooling_schemes = [self.global_pooling_schemes]
        if self.global_pooling_schemes is not None:
            for s in self.global_pooling_schemes:
                assert s in _POOL, f"Global pooling scheme {s} not supported."
        if self.add_global_variables_after_pooling:  # :151-155
            assert self.global_pooling_schemes
        act = self.activation_layer
        if act is None or act.lower() == "relu":  # :161-168
            self.activation_layer = "relu"
        elif act.lower() == "gelu":
            self.activation_layer = "gelu"
        else:
            raise ValueError(f"Activation layer {act} not supported.")
        self.nb_global_variables = 5 + self.nb_inputs  # :176

    # -- layer shapes, dynedge.py:184-250 ---------------------------------- #
    @property
    def nb_conv_inputs(self) -> int:
        n = self.nb_inputs
        if not self.add_global_variables_after_pooling:
            n += self.nb_global_variables
        return n

    def mlp_index(self, j: int) -> int:
        """Index of the j-th Linear inside a conv / post Sequential."""
        return j * (3 if self.add_norm_layer else 2)

    def linear_shapes(self) -> Dict[str, Tuple[int, int]]:
        """{state_dict prefix: (out_features, in_features)} of every Linear."""
        shapes: Dict[str, Tuple[int, int]] = {}
        lat = self.nb_conv_inputs
        for i, sizes in enumerate(self.dynedge_layer_sizes):
            dims = [lat] + list(sizes)
            for j, (a, b) in enumerate(zip(dims[:-1], dims[1:])):
                if j == 0:
                    a *= 2
                shapes[f"_conv_layers.{i}.nn.{self.mlp_index(j)}"] = (b, a)
            lat = dims[-1]
        lat = sum(s[-1] for s in self.dynedge_layer_sizes) + self.nb_conv_inputs
        dims = [lat] + list(self.post_processing_layer_sizes)
        for j, (a, b) in enumerate(zip(dims[:-1], dims[1:])):
            shapes[f"_post_processing.{self.mlp_index(j)}"] = (b, a)
        npool = len(self.global_pooling_schemes) if self.global_pooling_schemes else 1
        lat = dims[-1] * npool
        if self.add_global_variables_after_pooling:
            lat += self.nb_global_variables
        dims = [lat] + list(self.readout_layer_sizes)
        for j, (a, b) in enumerate(zip(dims[:-1], dims[1:])):
            shapes[f"_readout.{2 * j}"] = (b, a)  # readout never gets LayerNorm
        return shapes

    @property
    def nb_outputs(self) -> int:
        return self.readout_layer_sizes[-1]


def init_dynedge_state_dict(spec: DynEdgeSpec, seed: int = 0) -> Dict[str, Tensor]:
    """Random parameters with the reference's names and nn.Linear's default init.

    Constructs real `torch.nn.Linear` / `LayerNorm` modules in the order the
    reference does (dynedge.py:184-250) under `torch.manual_seed(seed)`, so a
    reference DynEdge built under the same seed holds identical values.
    """
    torch.manual_seed(seed)
    sd: Dict[str, Tensor] = {}
    for name, (out_f, in_f) in spec.linear_shapes().items():
        lin = torch.nn.Linear(in_f, out_f)
        sd[name + ".weight"] = lin.weight.detach().clone()
        sd[name + ".bias"] = lin.bias.detach().clone()
        if spec.add_norm_layer and not name.startswith("_readout"):
            head, idx = name.rsplit(".", 1)
            ln = f"{head}.{int(idx) + 1}"
            sd[ln + ".weight"] = torch.ones(out_f)
            sd[ln + ".bias"] = torch.zeros(out_f)
    return sd


def _act(spec: DynEdgeSpec, x: Tensor) -> Tensor:
    if spec.activation_layer == "relu":
        return torch.relu(x)
    return torch.nn.functional.gelu(x)


def _mlp(spec: DynEdgeSpec, sd: Dict[str, Tensor], prefix: str, n_lin: int,
         x: Tensor, norm: bool) -> Tensor:
    step = 3 if norm else 2
    for j in range(n_lin):
        p = f"{prefix}.{j * step}"
        x = torch.nn.functional.linear(x, sd[p + ".weight"], sd[p + ".bias"])
        if norm:
            q = f"{prefix}.{j * step + 1}"
            x = torch.nn.functional.layer_norm(
                x, (x.shape[-1],), sd[q + ".weight"], sd[q + ".bias"]
            )
        x = _act(spec, x)
    return x


def global_variables(x: Tensor, edge_index: Tensor, batch: Tensor,
                     n_pulses: Tensor) -> Tensor:
    """dynedge.py:267-294 with `torch.log10(data.n_pulses)` (:303-308)."""
    M = int(batch.max()) + 1
    h = [ops.homophily(edge_index, x[:, c], batch).reshape(-1, 1) for c in range(4)]
    means = ops.scatter_mean(x, batch, M)
    logn = torch.log10(n_pulses).to(x.dtype)
    return torch.cat([means] + h + [logn.unsqueeze(1)], dim=1)


def dynedge_forward(
    spec: DynEdgeSpec,
    sd: Dict[str, Tensor],
    x: Tensor,
    edge_index: Tensor,
    batch: Tensor,
    n_pulses: Tensor,
    dense_distribute: bool = False,
    return_intermediates: bool = False,
):
    """`DynEdge.forward(data)` (dynedge.py:296-350) on CPU tensors.

    `dense_distribute=True` executes the reference's literal [N,B] one-hot
    product (dynedge.py:309-320); the default uses `global_variables[batch]`,
    which is the same value (one non-zero term per row) without the N*B*G
    temporary.  The timed CPU baseline uses the literal form.
    """
    x = x.detach().cpu()
    edge_index = edge_index.detach().cpu().long()
    batch = batch.detach().cpu().long()
    n_pulses = n_pulses.detach().cpu()
    inter: Dict[str, Tensor] = {}

    gv = global_variables(x, edge_index, batch, n_pulses)
    inter["global_variables"] = gv

    if not spec.add_global_variables_after_pooling:
        if dense_distribute:
            distribute = (
                batch.unsqueeze(1) == torch.unique(batch).unsqueeze(0)
            ).type(torch.float)
            gvd = torch.sum(distribute.unsqueeze(2) * gv.unsqueeze(0), dim=1)
        else:
            # torch.unique(batch) enumerates the event ids that occur, in order
            uniq, inv = torch.unique(batch, return_inverse=True)
            gvd = gv[uniq][inv]
        x = torch.cat((x, gvd.to(x.dtype)), dim=1)

    skip = [x]
    graphs: List[Tensor] = [edge_index]
    for i, sizes in enumerate(spec.dynedge_layer_sizes):
        prefix = f"_conv_layers.{i}.nn"

        def nn_(t: Tensor, prefix=prefix, n=len(sizes)) -> Tensor:
            return _mlp(spec, sd, prefix, n, t, spec.add_norm_layer)

        x = ops.edge_conv(x, edge_index, nn_, aggr="add")  # layers.py:67
        edge_index = ops.knn_graph(  # layers.py:70-74
            x[:, spec.features_subset], spec.nb_neighbours, batch
        )
        skip.append(x)
        graphs.append(edge_index)
        inter[f"conv{i}"] = x
    inter["graphs"] = graphs  # graphs[i] = edge_index consumed by conv i

    x = torch.cat(skip, dim=1)
    x = _mlp(spec, sd, "_post_processing", len(spec.post_processing_layer_sizes),
             x, spec.add_norm_layer)
    inter["post"] = x

    if not spec.skip_readout:
        if spec.global_pooling_schemes:
            M = int(batch.max()) + 1
            x = torch.cat(
                [_POOL[s](x, batch, M) for s in spec.global_pooling_schemes], dim=1
            )
            inter["pooled"] = x
            if spec.add_global_variables_after_pooling:
                x = torch.cat([x, gv], dim=1)
        x = _mlp(spec, sd, "_readout", len(spec.readout_layer_sizes), x, False)

    if return_intermediates:
        return x, inter
    return x
