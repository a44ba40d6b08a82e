"""`DynEdgeJINST` on the CUDA kernels (fp32, or the 16-bit tensor-core kernels).

Mirrors /root/reference/src/graphnet/models/gnn/dynedge_jinst.py:19-160 - constructor arguments,
parameter names (`conv_add{1..4}.nn.{0,2}.*`, `nn1`, `nn2`, `nn3`) and output - so checkpoints of
the reference load.  It is DynEdge's older sibling: four DynEdgeConv layers with LeakyReLU MLPs
and "add" aggregation, the input and all four layer outputs concatenated, a two-layer node MLP,
max/min/sum/mean pooling with the xyzt homophily of the input graph and the raw pulse count
appended, and one output layer.  `precision="fp32"` (default) runs on the kernels of `DynEdge`'s
FFMA path; `precision="bf16"` / `"fp16"` on the tcgen05 kernels (node GEMMs, the fused edge kernel
with its LeakyReLU code, the node MLP), 2e-2 of the fp32 result.
"""

from __future__ import annotations

from typing import Any

import torch
from torch import Tensor

from . import ops
from .dynedge import DynEdgeConv


class DynEdgeJINST(torch.nn.Module):
    def __init__(self, nb_inputs: int, layer_size_scale: int = 4, precision: str = "fp32"):
        super().__init__()
        if precision not in ("fp32", "bf16", "fp16"):
            raise ValueError(f"precision {precision!r} not supported (fp32 | bf16 | fp16)")
        self.precision = precision
        c = layer_size_scale
        l1, l2, l3, l4, l5, l6 = nb_inputs, c * 16 * 2, c * 32 * 2, c * 42 * 2, c * 32 * 2, c * 16 * 2
        self._nb_inputs, self._nb_outputs = nb_inputs, l6
        self._dims = (l1, l3, l5)

        def conv(i: int, o: int, m: int) -> DynEdgeConv:
            return DynEdgeConv(
                torch.nn.Sequential(torch.nn.Linear(i * 2, m), torch.nn.LeakyReLU(),
                                    torch.nn.Linear(m, o), torch.nn.LeakyReLU()),
                aggr="add", nb_neighbors=8, features_subset=slice(0, 3))

        self.conv_add1 = conv(l1, l3, l2)
        self.conv_add2 = conv(l3, l3, l4)
        self.conv_add3 = conv(l3, l3, l4)
        self.conv_add4 = conv(l3, l3, l4)
        self.nn1 = torch.nn.Linear(l3 * 4 + l1, l4)
        self.nn2 = torch.nn.Linear(l4, l5)
        self.nn3 = torch.nn.Linear(4 * l5 + 5, l6)
        self.lrelu = torch.nn.LeakyReLU()

    @property
    def nb_inputs(self) -> int:
        return self._nb_inputs

    @property
    def nb_outputs(self) -> int:
        return self._nb_outputs

    def forward(self, data: Any) -> Tensor:
        if (self.training and torch.is_grad_enabled()
                and any(p.requires_grad for p in self.parameters())):
            # inference only: a silent forward without grad_fn would leave the backbone frozen
            # while an optimizer updates the task heads
            raise NotImplementedError(
                "graphnet_b200.DynEdgeJINST has no backward pass: call model.eval() or wrap the "
                "call in torch.no_grad()")
        with torch.no_grad():
            return self._forward(data) if self.precision == "fp32" else self._forward_16(data)

    def _forward_16(self, data: Any) -> Tensor:
        """The same forward on the tensor-core kernels (see dynedge_bf16.forward_bf16)."""
        from .dynedge import _WeightCache
        from .dynedge_bf16 import _conv_pack, _round_up, _slots
        x = data.x
        if not x.is_cuda:
            raise RuntimeError("graphnet_b200.DynEdgeJINST needs CUDA tensors (no CPU fallback)")
        x = x.float().contiguous()
        dev = x.device
        io = torch.float16 if self.precision == "fp16" else torch.bfloat16
        N, F = x.shape
        l1, l3, l5 = self._dims
        n_pulses = data.n_pulses.reshape(-1)
        B = int(n_pulses.shape[0])
        ptr, bflags = ops.batch_to_ptr(data.batch, B)
        graph = ops.coo_to_csr(data.edge_index, N)
        flags = graph.flags | bflags
        gv = ops.global_variables(x, graph, ptr, n_pulses)
        # skip buffer [x | conv1..4] with 8-element aligned slices, in the 16-bit type
        w0, w3 = _round_up(l1, 8), _round_up(l3, 8)
        total = w0 + 4 * w3
        skip = torch.zeros((N, total), dtype=io, device=dev)
        skip[:, :l1] = x.to(io)
        # the raw features enter the first node GEMM as bf16 [hi | lo | hi] (16 mantissa bits)
        x3 = ops.split_bf16(x, w0, terms=3)
        off_in, d_in, off_out = 0, l1, w0
        convs = [self.conv_add1, self.conv_add2, self.conv_add3, self.conv_add4]
        for li, conv in enumerate(convs):
            pk = _conv_pack(conv, io, first=(li == 0))
            PQ = torch.empty((N, 2 * pk.Kp), dtype=torch.float16, device=dev)
            if li == 0:
                ops.tc_linear_bf16(x3, pk.split_weights(w0), pk.bias_pq, PQ, K=3 * w0, act="none")
            else:
                ops.tc_linear_bf16(skip[:, off_in:], pk.Wpq, pk.bias_pq, PQ, K=d_in, act="none")
            last = li + 1 == len(convs)
            coords = None if last else torch.empty((N, 3), dtype=torch.float32, device=dev)
            ops.tc_edge_mlp_aggregate(PQ[:, :pk.Kp], PQ[:, pk.Kp:], pk.Kp, graph,
                                      _slots(conv.nb_neighbors), pk.W2p, pk.b, skip[:, off_out:],
                                      coords, [0, 1, 2], act=pk.act, aggr="add")
            if not last:      # the graph built after the last layer is never used
                graph = ops.knn_graph_ell(coords, conv.nb_neighbors, ptr, [0, 1, 2])
            off_in, d_in = off_out, l3
            off_out += w3
        cache = self.__dict__.setdefault("_cache_nn16", {}).setdefault(io, _WeightCache())

        def build():
            W1 = torch.zeros((self.nn1.out_features, total), dtype=torch.float32, device=dev)
            W = self.nn1.weight.detach().float()      # columns: [x (l1) | conv1..4 (l3 each)]
            W1[:, :l1] = W[:, :l1]
            for i in range(4):
                W1[:, w0 + i * w3:w0 + i * w3 + l3] = W[:, l1 + i * l3:l1 + (i + 1) * l3]
            k2 = _round_up(self.nn2.in_features, 8)
            W2 = torch.zeros((self.nn2.out_features, k2), dtype=torch.float32, device=dev)
            W2[:, :self.nn2.in_features] = self.nn2.weight.detach().float()
            return W1.to(io).contiguous(), W2.to(io).contiguous()
        W1, W2 = cache.get([self.nn1.weight, self.nn2.weight], build)
        h1 = torch.zeros((N, W2.shape[1]), dtype=io, device=dev) if W2.shape[1] != self.nn1.out_features \
            else torch.empty((N, W2.shape[1]), dtype=io, device=dev)
        ops.tc_linear_bf16(skip, W1, self.nn1.bias.detach().float(), h1, K=total, act="leaky_relu")
        h = torch.empty((N, l5), dtype=torch.float32, device=dev)
        ops.tc_linear_bf16(h1, W2, self.nn2.bias.detach().float(), h, K=self.nn2.in_features,
                           act="none")
        pooled = torch.empty((B, 4 * l5 + 5), dtype=torch.float32, device=dev)
        ops.segment_pool(h, ptr, ["max", "min", "sum", "mean"], out=pooled)
        pooled[:, 4 * l5] = gv[:, F + 3]          # h_t
        pooled[:, 4 * l5 + 1:4 * l5 + 4] = gv[:, F:F + 3]   # h_x, h_y, h_z
        pooled[:, 4 * l5 + 4] = n_pulses.to(torch.float32)
        torch.nn.functional.leaky_relu_(pooled, 0.01)
        out = ops.linear_f32(pooled, self.nn3.weight, self.nn3.bias, act="leaky_relu")
        if self.validate_inputs:
            from .dynedge import DynEdge
            DynEdge._check_flags(int(flags.item()))
        return out

    def _forward(self, data: Any) -> Tensor:
        x = data.x
        if not x.is_cuda:
            raise RuntimeError("graphnet_b200.DynEdgeJINST needs CUDA tensors (no CPU fallback)")
        x = x.float().contiguous()
        dev = x.device
        N, F = x.shape
        l1, l3, l5 = self._dims
        n_pulses = data.n_pulses.reshape(-1)
        B = int(n_pulses.shape[0])
        ptr, bflags = ops.batch_to_ptr(data.batch, B)
        graph = ops.coo_to_csr(data.edge_index, N)
        flags = graph.flags | bflags
        # homophily of x, y, z, t over the input graph (models/utils.py:17-33): columns F..F+3
        gv = ops.global_variables(x, graph, ptr, n_pulses)
        skip = torch.empty((N, l1 + 4 * l3), dtype=torch.float32, device=dev)
        skip[:, :l1] = x
        off_in, d_in, off_out = 0, l1, l1
        convs = [self.conv_add1, self.conv_add2, self.conv_add3, self.conv_add4]
        for li, conv in enumerate(convs):
            conv._conv_f32(skip[:, off_in:off_in + d_in], graph, out=skip[:, off_out:off_out + l3])
            if li + 1 < len(convs):      # the graph built after the last layer is never used
                graph = ops.knn_graph_ell(skip, conv.nb_neighbors, ptr,
                                          [off_out, off_out + 1, off_out + 2])
            off_in, d_in = off_out, l3
            off_out += l3
        h = ops.linear_f32(skip, self.nn1.weight, self.nn1.bias, act="leaky_relu")
        h = ops.linear_f32(h, self.nn2.weight, self.nn2.bias, act="none")
        pooled = torch.empty((B, 4 * l5 + 5), dtype=torch.float32, device=dev)
        ops.segment_pool(h, ptr, ["max", "min", "sum", "mean"], out=pooled)
        pooled[:, 4 * l5] = gv[:, F + 3]          # h_t
        pooled[:, 4 * l5 + 1:4 * l5 + 4] = gv[:, F:F + 3]   # h_x, h_y, h_z
        pooled[:, 4 * l5 + 4] = n_pulses.to(torch.float32)
        torch.nn.functional.leaky_relu_(pooled, 0.01)
        out = ops.linear_f32(pooled, self.nn3.weight, self.nn3.bias, act="leaky_relu")
        if self.validate_inputs:
            # one 4-byte device->host read per forward (as DynEdge): unsorted `batch`, edge targets
            # out of order or ids out of range raise instead of producing garbage
            from .dynedge import DynEdge
            DynEdge._check_flags(int(flags.item()))
        return out

    validate_inputs: bool = True
