"""Drop-in `DynEdge` / `DynEdgeConv` running on the sm_100a kernels.

Mirrors the reference's constructor signatures, parameter names (so existing
`state_dict`s load) and `forward` contracts:

  DynEdge      /root/reference/src/graphnet/models/gnn/dynedge.py:22-350
  DynEdgeConv  /root/reference/src/graphnet/models/components/layers.py:27-76

The modules hold ordinary `torch.nn.Linear` parameters; `forward` hands their
storage to the C ABI (`graphnet_b200.ops`).  There is no PyTorch/CPU fallback
path: CPU tensors are refused.
"""

from __future__ import annotations

from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple, Union

import torch
from torch import LongTensor, Tensor

from . import ops
from .ops import Graph

GLOBAL_POOLINGS = ("min", "max", "sum", "mean")  # dynedge.py:14-19


def _act_name(module: torch.nn.Module) -> str:
    if isinstance(module, torch.nn.ReLU):
        return "relu"
    if isinstance(module, torch.nn.GELU):
        if getattr(module, "approximate", "none") != "none":
            raise NotImplementedError("GELU(approximate=...) is not supported")
        return "gelu"
    if isinstance(module, torch.nn.LeakyReLU):
        if abs(module.negative_slope - 0.01) > 1e-12:
            raise NotImplementedError("LeakyReLU with a slope other than 0.01 is not supported")
        return "leaky_relu"
    raise NotImplementedError(f"activation {type(module).__name__} is not supported")


def _parse_mlp(seq: torch.nn.Module) -> Tuple[List[torch.nn.Linear], str]:
    """Split Sequential(Linear, act[, Linear, act ...]) into its Linears + activation."""
    mods = list(seq) if isinstance(seq, torch.nn.Sequential) else [seq]  # keeps shared modules
    linears: List[torch.nn.Linear] = []
    act = None
    i = 0
    while i < len(mods):
        if not isinstance(mods[i], torch.nn.Linear):
            raise NotImplementedError(
                f"unsupported layer {type(mods[i]).__name__} in MLP; the fused kernels take "
                "Linear -> activation sequences (add_norm_layer=True is not built yet)")
        linears.append(mods[i])
        if i + 1 >= len(mods):
            raise NotImplementedError("every Linear must be followed by an activation")
        a = _act_name(mods[i + 1])
        if act is not None and a != act:
            raise NotImplementedError("mixed activations inside one MLP")
        act = a
        i += 2
    assert act is not None
    return linears, act


def _apply_act_rows(h: Tensor, act: str) -> Tensor:
    if act == "none":
        return h
    # activation alone (first EdgeConv Linear without a norm layer after it): LayerNorm kernel
    # is not involved; a Linear-free activation only occurs on this unfused path
    if act == "relu":
        return torch.relu_(h)
    if act == "leaky_relu":
        return torch.nn.functional.leaky_relu_(h, 0.01)
    return torch.nn.functional.gelu(h)


def _parse_mlp_norm(seq: torch.nn.Module):
    """Sequential(Linear[, LayerNorm], act, ...) -> [(Linear, LayerNorm | None)], activation."""
    mods = list(seq) if isinstance(seq, torch.nn.Sequential) else [seq]
    steps, act, i = [], None, 0
    while i < len(mods):
        if not isinstance(mods[i], torch.nn.Linear):
            raise NotImplementedError(f"unsupported layer {type(mods[i]).__name__} in MLP")
        lin, ln = mods[i], None
        i += 1
        if i < len(mods) and isinstance(mods[i], torch.nn.LayerNorm):
            ln = mods[i]
            i += 1
        if i >= len(mods):
            raise NotImplementedError("every Linear must be followed by an activation")
        a = _act_name(mods[i])
        if act is not None and a != act:
            raise NotImplementedError("mixed activations inside one MLP")
        act = a
        i += 1
        steps.append((lin, ln))
    return steps, act


def _mlp_rows_norm(h: Tensor, steps, act: str) -> Tensor:
    """Row-wise MLP with LayerNorm between Linear and activation (fp32 kernels)."""
    for lin, ln in steps:
        if ln is None:
            h = ops.linear_f32(h, lin.weight, lin.bias, act=act)
        else:
            h = ops.linear_f32(h, lin.weight, lin.bias, act="none")
            ops.layernorm_act_(h, ln.weight, ln.bias, ln.eps, act=act)
    return h


class _WeightCache:
    """Repacked weights, rebuilt when a parameter is modified in place or replaced."""

    def __init__(self) -> None:
        self._key: Optional[Tuple] = None
        self._val: Any = None

    def get(self, params: Sequence[Tensor], build: Callable[[], Any]) -> Any:
        key = tuple((p.data_ptr(), p._version, str(p.device)) for p in params)
        if key != self._key:
            self._val = build()
            self._key = key
        return self._val


class DynEdgeConv(torch.nn.Module):
    """Dynamical edge convolution layer (layers.py:27-76).

    `forward(x, edge_index, batch) -> (x, edge_index)`: EdgeConv with `nn` and
    `aggr`, then `knn_graph(x[:, features_subset], nb_neighbors, batch)`.
    """

    def __init__(
        self,
        nn: Callable,
        aggr: str = "max",
        nb_neighbors: int = 8,
        features_subset: Optional[Union[Sequence[int], slice]] = None,
        **kwargs: Any,
    ):
        super().__init__()
        if features_subset is None:
            features_subset = slice(None)  # layers.py:52-53: use all features
        assert isinstance(features_subset, (list, slice))
        if aggr not in ops.AGGR:
            raise ValueError(f"aggr={aggr!r} not supported")
        self.nn = nn
        self.aggr = aggr
        self.nb_neighbors = nb_neighbors
        self.features_subset = features_subset
        self._cache_f32 = _WeightCache()

    # -- internals ---------------------------------------------------------- #
    def _mlp(self) -> Tuple[List[torch.nn.Linear], str]:
        linears, act = _parse_mlp(self.nn)
        if len(linears) > 2:
            raise NotImplementedError(
                "DynEdgeConv MLPs with more than two Linear layers are not built yet")
        return linears, act

    def _has_norm(self) -> bool:
        return any(isinstance(m, torch.nn.LayerNorm) for m in self.nn)

    @property
    def out_features(self) -> int:
        return [m for m in self.nn if isinstance(m, torch.nn.Linear)][-1].out_features

    def _conv_f32_norm(self, x: Tensor, graph: Graph, out: Optional[Tensor] = None) -> Tensor:
        """EdgeConv with LayerNorm inside the message MLP (add_norm_layer=True,
        gnn/dynedge.py:202-203): the normalisation needs every edge's full feature row, so the
        per-slot rows are materialised - the reference's own route - around the dense kernels."""
        steps, act = _parse_mlp_norm(self.nn)
        (l1, n1), rest = steps[0], steps[1:]
        N = x.shape[0]
        if graph.ell_stride <= 0:   # user edge_index: CSR -> slot table (one host read)
            graph = ops.csr_to_ell(graph, max(int(graph.max_degree.item()), 1))
        Wpq, _ = self._cache_f32.get([l1.weight], lambda: ops.edgeconv_prepare_f32(l1.weight, None))
        a = l1.out_features
        bias = torch.zeros(2 * a, dtype=torch.float32, device=x.device)
        if l1.bias is not None:
            bias[:a] = l1.bias.detach().float()
        pq = ops.linear_f32(x, Wpq, bias, act="none")                # [N, 2a] = [P | Q]
        h = ops.edge_preact_f32(pq, a, graph)                        # [N*S, a]
        if n1 is not None:
            ops.layernorm_act_(h, n1.weight, n1.bias, n1.eps, act=act)
        else:
            h = _apply_act_rows(h, act)
        h = _mlp_rows_norm(h, rest, act)
        if out is None:
            out = torch.empty((N, h.shape[1]), dtype=torch.float32, device=x.device)
        if self.aggr in ("add", "sum"):
            return ops.slot_sum_f32(h, graph, out)
        return ops.slot_aggregate_f32(h, graph, out, self.aggr)

    def _conv_f32(self, x: Tensor, graph: Graph, out: Optional[Tensor] = None) -> Tensor:
        n_lin = sum(isinstance(m, torch.nn.Linear) for m in self.nn)
        if self._has_norm() or n_lin > 2:   # the fused kernel covers one or two Linear layers
            return self._conv_f32_norm(x, graph, out)
        linears, act = self._mlp()
        l1 = linears[0]
        l2 = linears[1] if len(linears) == 2 else None
        params = [l1.weight] + ([l2.weight] if l2 is not None else [])
        Wpq, W2T = self._cache_f32.get(
            params,
            lambda: ops.edgeconv_prepare_f32(l1.weight, None if l2 is None else l2.weight))
        return ops.edgeconv_f32(
            x, graph, Wpq, l1.bias, W2T, None if l2 is None else l2.bias,
            act=act, aggr=self.aggr, out=out)

    def forward(self, x: Tensor, edge_index: LongTensor, batch: Optional[Tensor] = None
                ) -> Tuple[Tensor, LongTensor]:
        N = x.shape[0]
        x = x.float().contiguous()
        graph = ops.coo_to_csr(edge_index, N)
        flags = int(graph.flags.item())
        if flags & _lib_flags.FLAG_INDEX_OUT_OF_RANGE:
            raise ValueError("edge_index holds node ids outside [0, N)")
        if flags & _lib_flags.FLAG_UNSORTED_TARGETS:
            order = torch.sort(edge_index[1], stable=True).indices
            graph = ops.coo_to_csr(edge_index[:, order], N)
        out = self._conv_f32(x, graph)
        cols = ops._columns(self.features_subset, out.shape[1])
        new_edge_index = ops.knn_graph(out[:, cols], self.nb_neighbors, batch)
        return out, new_edge_index


from . import _lib as _lib_flags  # noqa: E402  (flag constants)


class _DataView:
    """`data` with another edge_index (the sorted copy), same attributes otherwise."""

    def __init__(self, data: Any, edge_index: Tensor):
        self._data = data
        self.edge_index = edge_index

    def __getattr__(self, name: str) -> Any:
        return getattr(self._data, name)


def _with_edge_index(data: Any, edge_index: Tensor) -> Any:
    return _DataView(data, edge_index)


class DynEdge(torch.nn.Module):
    """DynEdge (dynamical edge convolutional) model (dynedge.py:22-350).

    Same constructor arguments, parameter names and output as the reference.
    `precision` is the one addition: "fp32" (default; FFMA kernels, outputs
    within 1e-4 relative of the reference), or "bf16" / "fp16" (tcgen05
    tensor-core kernels with bf16 / fp16 activations, 2e-2).
    """

    def __init__(
        self,
        nb_inputs: int,
        *,
        nb_neighbours: int = 8,
        features_subset: Optional[Union[List[int], slice]] = None,
        dynedge_layer_sizes: Optional[List[Tuple[int, ...]]] = None,
        post_processing_layer_sizes: Optional[List[int]] = None,
        readout_layer_sizes: Optional[List[int]] = None,
        global_pooling_schemes: Optional[Union[str, List[str]]] = None,
        add_global_variables_after_pooling: bool = False,
        activation_layer: Optional[str] = None,
        add_norm_layer: bool = False,
        skip_readout: bool = False,
        precision: str = "fp32",
    ):
        super().__init__()
        if features_subset is None:
            features_subset = slice(0, 3)
        if dynedge_layer_sizes is None:
            dynedge_layer_sizes = [(128, 256), (336, 256), (336, 256), (336, 256)]
        assert isinstance(dynedge_layer_sizes, list)
        assert len(dynedge_layer_sizes)
        assert all(isinstance(sizes, tuple) for sizes in dynedge_layer_sizes)
        assert all(len(sizes) > 0 for sizes in dynedge_layer_sizes)
        assert all(all(size > 0 for size in sizes) for sizes in dynedge_layer_sizes)
        self._dynedge_layer_sizes = dynedge_layer_sizes

        if post_processing_layer_sizes is None:
            post_processing_layer_sizes = [336, 256]
        assert isinstance(post_processing_layer_sizes, list)
        assert len(post_processing_layer_sizes)
        assert all(size > 0 for size in post_processing_layer_sizes)
        self._post_processing_layer_sizes = post_processing_layer_sizes

        if readout_layer_sizes is None:
            readout_layer_sizes = [128]
        assert isinstance(readout_layer_sizes, list)
        assert len(readout_layer_sizes)
        assert all(size > 0 for size in readout_layer_sizes)
        self._readout_layer_sizes = readout_layer_sizes

        if isinstance(global_pooling_schemes, str):
            global_pooling_schemes = [global_pooling_schemes]
        if isinstance(global_pooling_schemes, list):
            for pooling_scheme in global_pooling_schemes:
                assert (
                    pooling_scheme in GLOBAL_POOLINGS
                ), f"Global pooling scheme {pooling_scheme} not supported."
        else:
            assert global_pooling_schemes is None
        self._global_pooling_schemes = global_pooling_schemes

        if add_global_variables_after_pooling:
            assert self._global_pooling_schemes, (
                "No global pooling schemes were request, so cannot add global"
                " variables after pooling."
            )
        self._add_global_variables_after_pooling = add_global_variables_after_pooling

        if activation_layer is None or activation_layer.lower() == "relu":
            activation: torch.nn.Module = torch.nn.ReLU()
        elif activation_layer.lower() == "gelu":
            activation = torch.nn.GELU()
        else:
            raise ValueError(f"Activation layer {activation_layer} not supported.")
        if precision not in ("fp32", "bf16", "fp16"):
            raise ValueError(f"precision {precision!r} not supported (fp32 | bf16 | fp16)")

        self._activation = activation
        self._nb_inputs = nb_inputs
        self._nb_outputs = self._readout_layer_sizes[-1]
        self._nb_global_variables = 5 + nb_inputs
        self._nb_neighbours = nb_neighbours
        self._features_subset = features_subset
        self._add_norm_layer = add_norm_layer
        self._skip_readout = skip_readout
        self.precision = precision
        self._construct_layers()

    # gnn/gnn.py:23-31
    @property
    def nb_inputs(self) -> int:
        return self._nb_inputs

    @property
    def nb_outputs(self) -> int:
        return self._nb_outputs

    def _construct_layers(self) -> None:
        """Same module tree as dynedge.py:184-250 (hence the same state_dict keys)."""
        nb_input_features = self._nb_inputs
        if not self._add_global_variables_after_pooling:
            nb_input_features += self._nb_global_variables
        self._nb_conv_inputs = nb_input_features

        self._conv_layers = torch.nn.ModuleList()
        nb_latent_features = nb_input_features
        for sizes in self._dynedge_layer_sizes:
            layers: List[torch.nn.Module] = []
            layer_sizes = [nb_latent_features] + list(sizes)
            for ix, (nb_in, nb_out) in enumerate(zip(layer_sizes[:-1], layer_sizes[1:])):
                if ix == 0:
                    nb_in *= 2
                layers.append(torch.nn.Linear(nb_in, nb_out))
                if self._add_norm_layer:
                    layers.append(torch.nn.LayerNorm(nb_out))
                layers.append(self._activation)
            self._conv_layers.append(
                DynEdgeConv(
                    torch.nn.Sequential(*layers),
                    aggr="add",
                    nb_neighbors=self._nb_neighbours,
                    features_subset=self._features_subset,
                )
            )
            nb_latent_features = nb_out

        nb_latent_features = (
            sum(sizes[-1] for sizes in self._dynedge_layer_sizes) + nb_input_features
        )
        post_processing_layers: List[torch.nn.Module] = []
        layer_sizes = [nb_latent_features] + list(self._post_processing_layer_sizes)
        for nb_in, nb_out in zip(layer_sizes[:-1], layer_sizes[1:]):
            post_processing_layers.append(torch.nn.Linear(nb_in, nb_out))
            if self._add_norm_layer:
                post_processing_layers.append(torch.nn.LayerNorm(nb_out))
            post_processing_layers.append(self._activation)
        self._post_processing = torch.nn.Sequential(*post_processing_layers)

        nb_poolings = len(self._global_pooling_schemes) if self._global_pooling_schemes else 1
        nb_latent_features = nb_out * nb_poolings
        if self._add_global_variables_after_pooling:
            nb_latent_features += self._nb_global_variables
        readout_layers: List[torch.nn.Module] = []
        layer_sizes = [nb_latent_features] + list(self._readout_layer_sizes)
        for nb_in, nb_out in zip(layer_sizes[:-1], layer_sizes[1:]):
            readout_layers.append(torch.nn.Linear(nb_in, nb_out))
            readout_layers.append(self._activation)
        self._readout = torch.nn.Sequential(*readout_layers)

    # ------------------------------------------------------------------ forward
    def forward(self, data: Any) -> Tensor:
        """Apply the forward pass (dynedge.py:296-350) to a collated batch.

        `data` needs `x` [N,F] float32, `edge_index` [2,E] int64, `batch` [N]
        int64 sorted, `n_pulses` [B] (any object with those attributes, e.g. a
        torch_geometric `Batch`).
        """
        if self._add_norm_layer and self.precision != "fp32":
            raise NotImplementedError("add_norm_layer=True runs on precision='fp32'")
        wants_grad = (self.training and torch.is_grad_enabled()
                      and any(p.requires_grad for p in self.parameters()))
        if wants_grad and self.precision != "bf16":
            # the hand-written backward exists for the bf16 kernels only; returning an output
            # without grad_fn here would let an optimizer step update the task heads while the
            # backbone silently stays frozen
            raise NotImplementedError(
                f"DynEdge(precision={self.precision!r}) has no backward pass: train with "
                "precision='bf16', or call model.eval() / wrap the call in torch.no_grad() for "
                "inference")
        if not wants_grad:   # inference kernels allocate nothing for autograd
            with torch.no_grad():
                return self._forward_checked(data)
        return self._forward_checked(data)

    def _forward_checked(self, data: Any) -> Tensor:
        out = self._dispatch(data, None)
        if self.validate_inputs and self._last_flags is not None:
            # one 4-byte device->host read per forward; set validate_inputs=False to
            # keep the call fully asynchronous when the inputs are known to be sane
            f = int(self._last_flags.item())
            if f & _lib_flags.FLAG_UNSORTED_TARGETS and not (
                    f & (_lib_flags.FLAG_INDEX_OUT_OF_RANGE | _lib_flags.FLAG_UNSORTED_BATCH)):
                # the reference accepts edges in any order: group them by target
                # (stable, keeps the per-target edge order) and run again
                order = torch.sort(data.edge_index[1], stable=True).indices
                out = self._dispatch(data, data.edge_index[:, order])
                f = int(self._last_flags.item())
            self._check_flags(f)
        return out

    validate_inputs: bool = True
    # 16-bit inference: enqueue the whole forward from ONE C-ABI call (csrc/forward.cu) where the
    # model's shape allows; False keeps the call-by-call Python sequence (same kernels, same bits)
    native_driver: bool = True
    # precision="fp32": multiply on the tensor cores with operands split into 16-bit pairs
    # (dynedge_x3.py; fp32 inputs, weights, accumulation and outputs).  False: the FFMA kernels,
    # the parity anchor (plain fp32 arithmetic in the reference's order of operations)
    fp32_tensor_cores: bool = True
    # 16-bit inference: run the edge kernel with k slots per node instead of k + 1 (see
    # dynedge_bf16._slots_min: pays off only where coincident kNN coordinates are rare)
    min_edge_slots: bool = False
    _last_flags: Optional[Tensor] = None
    # k of the input graph built on the device when a batch carries no `edge_index`
    # (None: the model's nb_neighbours, the pairing every reference example uses)
    input_graph_neighbours: Optional[int] = None

    def _dispatch(self, data: Any, edge_index: Optional[Tensor]) -> Tensor:
        if (self.training and self.precision == "bf16" and torch.is_grad_enabled()
                and any(p.requires_grad for p in self.parameters())):
            # train mode with autograd recording: the forward that keeps what the hand-written
            # backward needs (fp32 / fp16 never get here: `forward` raises for them)
            from .training import dynedge_train_forward
            if edge_index is not None:
                data = _with_edge_index(data, edge_index)
            return dynedge_train_forward(self, data)
        if self.precision in ("bf16", "fp16"):
            from .dynedge_bf16 import forward_bf16
            return forward_bf16(self, data, edge_index)
        if self.fp32_tensor_cores:
            from .dynedge_x3 import eligible, forward_f32_tc
            if eligible(self):
                return forward_f32_tc(self, data, edge_index)
        return self._forward_f32(data, edge_index=edge_index)

    def _graph_inputs(self, data: Any, edge_index: Optional[Tensor] = None
                      ) -> Tuple[Tensor, Tensor, Tensor, Graph, int, int]:
        x = data.x
        if not x.is_cuda:
            raise RuntimeError("graphnet_b200.DynEdge needs CUDA tensors (no CPU fallback)")
        x = x.float().contiguous()
        N = x.shape[0]
        n_pulses = data.n_pulses
        B = int(n_pulses.shape[0]) if n_pulses.dim() > 0 else 1
        n_pulses = n_pulses.reshape(-1)
        ptr, bflags = ops.batch_to_ptr(data.batch, B)
        if edge_index is None:
            edge_index = getattr(data, "edge_index", None)
        if edge_index is None:
            # edges on the device: a batch that arrives without `edge_index`
            # gets the graph the reference's KNNGraph would have built per event on the host
            # (graphs/graphs.py:24-73: columns [0, 1, 2], k = nb_nearest_neighbours) from ONE
            # launch, kept in the slot-table layout the kernels consume - no COO, no host sync
            k = self.input_graph_neighbours or self._nb_neighbours
            graph = ops.knn_graph_ell(x, k, ptr, [0, 1, 2])
            graph.flags = bflags
        else:
            graph = ops.coo_to_csr(edge_index, N)
            graph.flags |= bflags
        return x, n_pulses, ptr, graph, N, B

    @staticmethod
    def _check_flags(f: int) -> None:
        if f & _lib_flags.FLAG_INDEX_OUT_OF_RANGE:
            raise ValueError("edge_index / batch hold ids out of range")
        if f & _lib_flags.FLAG_UNSORTED_BATCH:
            raise ValueError("`batch` must be sorted (PyG batching invariant)")
        if f & _lib_flags.FLAG_UNSORTED_TARGETS:
            raise ValueError(
                "edge_index targets (row 1) must be non-decreasing, as knn_graph and "
                "Batch.from_data_list produce them")

    def _forward_f32(self, data: Any, return_intermediates: bool = False,
                     edge_index: Optional[Tensor] = None):
        x, n_pulses, ptr, graph, N, B = self._graph_inputs(data, edge_index)
        self._last_flags = graph.flags
        dev = x.device
        F = self._nb_inputs
        assert x.shape[1] == F, f"expected {F} input features, got {x.shape[1]}"
        act = _act_name(self._activation)
        conv_out = [conv.out_features for conv in self._conv_layers]
        D0 = self._nb_conv_inputs
        width = D0 + sum(conv_out)
        # skip-cat buffer: [x0 | conv1 | conv2 | ...]; every producer writes its slice
        skip = torch.empty((N, width), dtype=torch.float32, device=dev)
        if self._add_global_variables_after_pooling:
            gv = ops.global_variables(x, graph, ptr, n_pulses)
            skip[:, :F] = x
        else:
            gv = ops.global_variables(x, graph, ptr, n_pulses, x0=skip, x0_cols=D0)
        inter: Dict[str, Any] = {"global_variables": gv, "graphs": []}

        off_in, d_in = 0, D0
        off_out = D0
        for li, conv in enumerate(self._conv_layers):
            xin = skip[:, off_in:off_in + d_in]
            xout = skip[:, off_out:off_out + conv_out[li]]
            conv._conv_f32(xin, graph, out=xout)
            if li + 1 < len(self._conv_layers) or return_intermediates:
                cols = [off_out + c for c in ops._columns(conv.features_subset, conv_out[li])]
                graph = ops.knn_graph_ell(skip, conv.nb_neighbors, ptr, cols)
                if return_intermediates:
                    inter["graphs"].append(graph)
            off_in, d_in = off_out, conv_out[li]
            off_out += conv_out[li]

        post, pact = _parse_mlp_norm(self._post_processing)
        h = _mlp_rows_norm(skip, post, pact)
        inter["post"] = h

        if not self._skip_readout:
            if self._global_pooling_schemes:
                P = len(self._global_pooling_schemes)
                C_ = h.shape[1]
                extra = self._nb_global_variables if self._add_global_variables_after_pooling else 0
                pooled = torch.empty((B, C_ * P + extra), dtype=torch.float32, device=dev)
                ops.segment_pool(h, ptr, self._global_pooling_schemes, out=pooled)
                if extra:
                    pooled[:, C_ * P:] = gv
                inter["pooled"] = pooled
                h = pooled
            ro, ract = _parse_mlp(self._readout)
            for lin in ro:
                h = ops.linear_f32(h, lin.weight, lin.bias, act=ract)
        if return_intermediates:
            inter["skip"] = skip
            return h, inter
        return h

    def check_inputs(self, data: Any) -> None:
        """Validate `batch` / `edge_index` ordering on the device (one host sync)."""
        _, _, _, graph, _, _ = self._graph_inputs(data)
        self._check_flags(int(graph.flags.item()))
