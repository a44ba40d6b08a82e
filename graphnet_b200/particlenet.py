"""`ParticleNeT` on the CUDA kernels (fp32, inference).

Mirrors /root/reference/src/graphnet/models/gnn/particlenet.py:22-256 - constructor arguments,
parameter names (`_conv_layers.{i}.nn.{j}.*`, `_readout.{j}.*`) and output - so checkpoints of the
reference load.  It is DynEdge's other sibling: DynEdgeConv blocks whose MLPs are
Linear - BatchNorm1d - activation x 3 with aggr="mean", the graph rebuilt after every block
(`dynamic=True`) or kept (`dynamic=False`), no skip concatenation, global pooling and a readout
MLP with Dropout.

In eval mode BatchNorm1d is the per-channel affine map of its running statistics, which is folded
into the Linear before it (W' = s W, b' = s (b - mean) + beta, s = gamma / sqrt(var + eps)); Dropout
is the identity.  The folded layers run on the same kernels as DynEdge's fp32 path.  Train mode
(batch statistics over all edges, dropout masks) is not built: `forward` raises.
"""

from __future__ import annotations

from typing import Any, List, Optional, Tuple, Union

import torch
from torch import Tensor

from . import ops
from .dynedge import DynEdgeConv, _WeightCache, _act_name

GLOBAL_POOLINGS = ("min", "max", "sum", "mean")


class ParticleNeT(torch.nn.Module):
    def __init__(
        self,
        nb_inputs: int,
        *,
        nb_neighbours: int = 16,
        features_subset: Optional[Union[List[int], slice]] = None,
        dynamic: bool = True,
        dynedge_layer_sizes: Optional[List[Tuple[int, ...]]] = None,
        readout_layer_sizes: Optional[List[int]] = None,
        global_pooling_schemes: Optional[Union[str, List[str]]] = "mean",
        activation_layer: Optional[str] = "relu",
        add_batchnorm_layer: bool = True,
        dropout_readout: float = 0.1,
        skip_readout: bool = False,
    ):
        super().__init__()
        if features_subset is None:
            features_subset = slice(0, 3)
        if dynedge_layer_sizes is None:
            dynedge_layer_sizes = [(64, 64, 64), (128, 128, 128), (256, 256, 256)]
        dynedge_layer_sizes = [tuple(s) if isinstance(s, list) else s for s in dynedge_layer_sizes]
        assert len(dynedge_layer_sizes)
        assert all(isinstance(s, tuple) and len(s) > 0 and all(v > 0 for v in s)
                   for s in dynedge_layer_sizes)
        if readout_layer_sizes is None:
            readout_layer_sizes = [256]
        assert isinstance(readout_layer_sizes, list) and len(readout_layer_sizes)
        assert all(size > 0 for size in readout_layer_sizes)
        if isinstance(global_pooling_schemes, str):
            global_pooling_schemes = [global_pooling_schemes]
        if isinstance(global_pooling_schemes, list):
            for scheme in global_pooling_schemes:
                assert scheme in GLOBAL_POOLINGS, f"Global pooling scheme {scheme} not supported."
        else:
            assert global_pooling_schemes is None
        if activation_layer is None or activation_layer.lower() == "relu":
            act: torch.nn.Module = torch.nn.ReLU()
        elif activation_layer.lower() == "gelu":
            act = torch.nn.GELU()
        else:
            raise ValueError(f"Activation layer {activation_layer} not supported.")
        self._activation = act
        self._nb_inputs = nb_inputs
        self._nb_outputs = readout_layer_sizes[-1]
        self._nb_neighbours = nb_neighbours
        self._features_subset = features_subset
        self._dynamic = dynamic
        self._dynedge_layer_sizes = dynedge_layer_sizes
        self._readout_layer_sizes = readout_layer_sizes
        self._global_pooling_schemes = global_pooling_schemes
        self._add_batchnorm_layer = add_batchnorm_layer
        self._dropout_readout = dropout_readout
        self._skip_readout = skip_readout

        # layers in the reference's construction order (particlenet.py:176-222)
        self._conv_layers = torch.nn.ModuleList()
        nb_latent = nb_inputs
        for sizes in dynedge_layer_sizes:
            layers: List[torch.nn.Module] = []
            dims = [nb_latent] + list(sizes)
            for ix, (nb_in, nb_out) in enumerate(zip(dims[:-1], dims[1:])):
                if ix == 0:
                    nb_in *= 2
                layers.append(torch.nn.Linear(nb_in, nb_out))
                if add_batchnorm_layer:
                    layers.append(torch.nn.BatchNorm1d(nb_out))
                layers.append(act)
            block = torch.nn.Module()
            block.nn = torch.nn.Sequential(*layers)
            self._conv_layers.append(block)
            nb_latent = dims[-1]
        nb_pool = len(global_pooling_schemes) if global_pooling_schemes else 1
        dims = [nb_latent * nb_pool] + list(readout_layer_sizes)
        readout: List[torch.nn.Module] = []
        for nb_in, nb_out in zip(dims[:-1], dims[1:]):
            readout += [torch.nn.Linear(nb_in, nb_out), act, torch.nn.Dropout(dropout_readout)]
        self._readout = torch.nn.Sequential(*readout)
        self._folded = [_WeightCache() for _ in dynedge_layer_sizes]

    @property
    def nb_inputs(self) -> int:
        return self._nb_inputs

    @property
    def nb_outputs(self) -> int:
        return self._nb_outputs

    def _folded_conv(self, i: int) -> DynEdgeConv:
        """DynEdgeConv over Linear-activation layers with the BatchNorm maps folded in."""
        seq = self._conv_layers[i].nn
        mods = list(seq)
        tensors = [t for m in mods for t in list(m.parameters(recurse=False)) +
                   list(m.buffers(recurse=False))]

        def build() -> DynEdgeConv:
            layers: List[torch.nn.Module] = []
            j = 0
            while j < len(mods):
                lin = mods[j]
                assert isinstance(lin, torch.nn.Linear)
                W = lin.weight.detach().float()
                b = (lin.bias.detach().float() if lin.bias is not None
                     else torch.zeros(W.shape[0], device=W.device))
                j += 1
                if j < len(mods) and isinstance(mods[j], torch.nn.BatchNorm1d):
                    bn = mods[j]
                    s = torch.rsqrt(bn.running_var.float() + bn.eps)
                    if bn.weight is not None:
                        s = s * bn.weight.detach().float()
                    b = (b - bn.running_mean.float()) * s
                    if bn.bias is not None:
                        b = b + bn.bias.detach().float()
                    W = W * s[:, None]
                    j += 1
                f = torch.nn.Linear(W.shape[1], W.shape[0]).to(W.device)
                f.weight.data.copy_(W)
                f.bias.data.copy_(b)
                f.requires_grad_(False)
                layers += [f, mods[j]]
                j += 1
            return DynEdgeConv(torch.nn.Sequential(*layers), aggr="mean",
                               nb_neighbors=self._nb_neighbours,
                               features_subset=self._features_subset)
        return self._folded[i].get(tensors, build)

    @torch.no_grad()
    def forward(self, data: Any) -> Tensor:
        x = data.x
        if not x.is_cuda:
            raise RuntimeError("graphnet_b200.ParticleNeT needs CUDA tensors (no CPU fallback)")
        if self.training:
            raise NotImplementedError(
                "ParticleNeT runs in eval mode here (BatchNorm running statistics folded into the "
                "Linear layers, Dropout off); call model.eval()")
        x = x.float().contiguous()
        N = x.shape[0]
        batch = data.batch
        B = int(data.n_pulses.reshape(-1).shape[0]) if hasattr(data, "n_pulses") and \
            data.n_pulses is not None else int(batch.max().item()) + 1
        ptr, _ = ops.batch_to_ptr(batch, B)
        graph = ops.coo_to_csr(data.edge_index, N)
        for i in range(len(self._conv_layers)):
            conv = self._folded_conv(i)
            x = conv._conv_f32(x, graph)
            if self._dynamic and i + 1 < len(self._conv_layers):   # the last rebuilt graph is unused
                cols = ops._columns(conv.features_subset, x.shape[1])
                graph = ops.knn_graph_ell(x, conv.nb_neighbors, ptr, cols)
        if self._skip_readout:
            return x
        if self._global_pooling_schemes:
            x = ops.segment_pool(x, ptr, self._global_pooling_schemes)
        for m in self._readout:
            if isinstance(m, torch.nn.Linear):
                x = ops.linear_f32(x, m.weight, m.bias, act=_act_name(self._activation))
        return x
