"""fp32 forward of DynEdge on the tensor cores (reference dynedge.py:296-350 at its own precision).

Every Linear keeps fp32 inputs, weights and outputs; what the tensor cores multiply are PAIRS of
16-bit matrices whose sum is the fp32 operand:

  * node-level Linear layers: the input is split into bf16 pieces x = x1 + x2 + x3 (gn_split_bf16),
    the weights likewise, and ONE bf16 GEMM over the concatenated pieces accumulates the product
    terms in fp32 (gn_tc_linear).  The factored first Linear of every EdgeConv MLP takes six terms,
    [x1 | x1 | x1 | x2 | x2 | x3] x [w1 | w2 | w3 | w1 | w2 | w1] - everything down to 2^-16 of the
    leading term, i.e. fp32-grade: its output decides the next layer's kNN graph, where an error
    of 1e-5 already flips near-ties.  The post-processing MLP takes three, [x1 | x2 | x1] x
    [w1 | w1 | w2] (16 bits in both factors; nothing downstream is discontinuous);
  * the edge kernel (gn_tc_edge_mlp_aggregate_f32): h = relu(P_i + Q_j) is formed in fp32 and fed
    as fp16 (hi, lo), W2 likewise, three MMAs per K step - 22 bits;
  * the readout runs on the fp32 FFMA kernel (a [B, 1024] x [1024, 128] product).

Outputs agree with the FFMA path (`DynEdge.fp32_tensor_cores = False`, the parity anchor) to a few
1e-6 relative and with the fp32 CPU oracle within the 1e-4 of BASELINE.json.
"""

from __future__ import annotations

from typing import Any, List, Optional

import torch
from torch import Tensor

from . import ops
from .dynedge import _parse_mlp, _WeightCache
from .dynedge_bf16 import _conv_pack, _round_up, _skip_layout, _slots


def _split3(W: Tensor, segments: List[tuple]) -> Tensor:
    """fp32 [n, sum(w)] -> bf16 [n, 3 * sum(kc)]: per (w, kc) segment [W_hi | W_hi | W_lo], each
    third kc wide (zero padded) - the partner of a [hi | lo | hi] input."""
    n = W.shape[0]
    total = sum(kc for _, kc in segments)
    out = torch.zeros((n, 3 * total), dtype=torch.bfloat16, device=W.device)
    src = dst = 0
    for w, kc in segments:
        Ws = W[:, src:src + w].float()
        hi = Ws.to(torch.bfloat16)
        lo = (Ws - hi.float()).to(torch.bfloat16)
        out[:, dst:dst + w] = hi
        out[:, dst + kc:dst + kc + w] = hi
        out[:, dst + 2 * kc:dst + 2 * kc + w] = lo
        src += w
        dst += 3 * kc
    return out.contiguous()


def _split6(W: Tensor, d: int, kc: int) -> Tensor:
    """fp32 [n, d] -> bf16 [n, 6 kc] = [w1 | w2 | w3 | w1 | w2 | w1] (pieces of W, blocks kc wide):
    the partner of a [x1 | x1 | x1 | x2 | x2 | x3] input."""
    W = W.float()
    w1 = W.to(torch.bfloat16)
    r1 = W - w1.float()
    w2 = r1.to(torch.bfloat16)
    w3 = (r1 - w2.float()).to(torch.bfloat16)
    out = torch.zeros((W.shape[0], 6 * kc), dtype=torch.bfloat16, device=W.device)
    for blk, piece in enumerate((w1, w2, w3, w1, w2, w1)):
        out[:, blk * kc:blk * kc + d] = piece
    return out.contiguous()


class _X3Pack:
    """Weights of a model in the layouts above (rebuilt when a parameter changes)."""

    def __init__(self, model):
        D0, widths, offs, total = _skip_layout(model)
        self.D0, self.widths, self.offs, self.total = D0, widths, offs, total
        self.convs = []
        d_in = D0
        for li, conv in enumerate(model._conv_layers):
            if conv.aggr not in ("add", "sum"):
                raise NotImplementedError("fp32 tensor-core path: aggr='add'")
            pk = _conv_pack(conv, torch.bfloat16, first=True)   # fp32 Wpq, fp16 W2p (+ b2 columns)
            if pk.act != "relu" or pk.Kp > 352:
                raise NotImplementedError("fp32 tensor-core path: ReLU, hidden <= 350")
            assert pk.d == d_in
            kc = _round_up(d_in, 8)
            linears, _ = conv._mlp()
            l2 = linears[1]
            W2 = torch.zeros((pk.b16, pk.Kp), dtype=torch.float32, device=l2.weight.device)
            W2[:pk.b, :pk.a] = l2.weight.detach().float()
            hi = pk.W2p                      # fp16(W2), b2 as (hi, lo) in the two "ones" columns
            lo = torch.zeros_like(hi)
            lo[:, :pk.a] = (W2[:, :pk.a] - hi[:, :pk.a].float()).to(torch.float16)
            cols = sorted(ops._columns(conv.features_subset, pk.b))
            self.convs.append({"pk": pk, "kc": kc, "Wpq6": _split6(pk.Wpq, d_in, kc),
                               "W2hi": hi.contiguous(), "W2lo": lo.contiguous(), "cols": cols,
                               "k": conv.nb_neighbors})
            d_in = pk.b
        linears, act = _parse_mlp(model._post_processing)
        if act != "relu":
            raise NotImplementedError("fp32 tensor-core path: ReLU post-processing")
        self.post = []
        segs = [(D0, offs[0])] + [(w, _round_up(w, 8)) for w in widths]
        for i, lin in enumerate(linears):
            W = lin.weight.detach().float()
            if i > 0:
                segs = [(W.shape[1], _round_up(W.shape[1], 8))]
            bias = None if lin.bias is None else lin.bias.detach().float().contiguous()
            self.post.append((_split3(W, segs), bias, W.shape[0], 3 * sum(kc for _, kc in segs)))


def _x3_pack(model) -> Optional[_X3Pack]:
    cache = model.__dict__.get("_cache_x3")
    if cache is None:
        cache = model.__dict__["_cache_x3"] = _WeightCache()

    def build():
        try:
            return _X3Pack(model)
        except NotImplementedError:
            return None
    return cache.get(list(model.parameters()), build)


def eligible(model) -> bool:
    """Models the tensor-core fp32 path covers; the others run on the FFMA kernels."""
    if model._add_norm_layer or model._add_global_variables_after_pooling:
        return False
    return _x3_pack(model) is not None


def forward_f32_tc(model, data: Any, edge_index: Optional[Tensor] = None) -> Tensor:
    pack = _x3_pack(model)
    assert pack is not None
    x, n_pulses, ptr, graph, N, B = model._graph_inputs(data, edge_index)
    model._last_flags = graph.flags
    dev = x.device
    F = model._nb_inputs
    assert x.shape[1] == F, f"expected {F} input features, got {x.shape[1]}"
    n_pulses = n_pulses.to(torch.int32).contiguous()   # (torch kernels stay outside the capture)
    batch = getattr(data, "batch", None)
    if isinstance(batch, Tensor):
        batch = batch.contiguous()

    def body() -> Tensor:
        # only this library's launches and torch.empty from here on: ops.run_captured replays
        # them as one CUDA graph (no idle GPU between the ~35 kernels)
        g = graph
        offs, total = pack.offs, pack.total
        x0w = offs[0]
        # S: every layer's features as bf16 [hi | lo | hi] triplets, [x0 | conv1 | conv2 | ...]; the
        # node GEMM of a layer reads its own triplet, the post-processing MLP the whole row
        S = torch.empty((N, 3 * total), dtype=torch.bfloat16, device=dev)
        xl = torch.empty((N, x0w), dtype=torch.float32, device=dev)     # layer input, fp32
        ops.global_variables(x, g, ptr, n_pulses, x0=S, x0_cols=x0w, x0b=xl, x0b_cols=x0w,
                             x0_split=True)
        A6 = ops.split_bf16(xl, x0w, terms=6)
        for li, cv in enumerate(pack.convs):
            pk, kc = cv["pk"], cv["kc"]
            PQ = torch.empty((N, 2 * pk.Kp), dtype=torch.float32, device=dev)
            ops.tc_linear_bf16(A6, cv["Wpq6"], pk.bias_pq, PQ, K=6 * kc, act="none")
            last = li + 1 == len(pack.convs)
            cols = cv["cols"]
            xl = torch.empty((N, pk.b), dtype=torch.float32, device=dev)
            coords = None if last else torch.empty((N, len(cols)), dtype=torch.float32, device=dev)
            ops.tc_edge_mlp_aggregate_f32(PQ[:, :pk.Kp], PQ[:, pk.Kp:], pk.Kp, g, _slots(cv["k"]),
                                          cv["W2hi"], cv["W2lo"], pk.b, xl, coords, cols)
            if last:
                ops.split_bf16(xl, _round_up(pk.b, 8), terms=3, out=S[:, 3 * offs[li]:])
            else:   # three terms for the post-processing MLP, six for the next node GEMM: one read
                A6 = ops.split_bf16_both(xl, _round_up(pk.b, 8), S[:, 3 * offs[li]:])
                g = ops.knn_graph_ell(coords, cv["k"], ptr, list(range(len(cols))))
        h16 = S
        pools = model._global_pooling_schemes if not model._skip_readout else None
        fuse_pool = bool(pools) and isinstance(batch, Tensor) and batch.dtype == torch.int64 \
            and pack.post[-1][2] % 2 == 0
        h = pooled = None
        for i, (Wp, bias, n, K) in enumerate(pack.post):
            lastp = i + 1 == len(pack.post)
            if lastp and fuse_pool:
                pooled = torch.empty((B, n * len(pools)), dtype=torch.float32, device=dev)
                ops.tc_linear_pool(h16, Wp, bias, K, batch, ptr, pools, pooled)
                break
            h = torch.empty((N, n), dtype=torch.float32, device=dev)
            ops.tc_linear_bf16(h16, Wp, bias, h, K=K, act="relu")
            if not lastp:
                h16 = ops.split_bf16(h, _round_up(n, 8), terms=3)
        if model._skip_readout:
            return h
        if pools and pooled is None:
            pooled = ops.segment_pool(h, ptr, pools)
        out = pooled if pooled is not None else h
        ro, ract = _parse_mlp(model._readout)
        for lin in ro:
            out = ops.linear_f32(out, lin.weight, lin.bias, act=ract)
        return out

    return ops.run_captured(id(pack), dev, body)


def time_edge_kernel_f32(model, data: Any, flush_buf: Tensor, layer: int = 1, reps: int = 5):
    """CUDA-event time of the fp32 edge kernel of conv `layer` alone (bench.py's fp32 rider)."""
    pack = _x3_pack(model)
    assert pack is not None
    x, n_pulses, ptr, graph, N, B = model._graph_inputs(data, None)
    dev = x.device
    cv = pack.convs[layer]
    pk = cv["pk"]
    g = torch.Generator(device=dev).manual_seed(0)
    PQ = torch.randn((N, 2 * pk.Kp), generator=g, device=dev) * 0.5
    out = torch.empty((N, pk.b), dtype=torch.float32, device=dev)
    coords = torch.empty((N, 3), dtype=torch.float32, device=dev)
    ell = ops.knn_graph_ell(x, cv["k"], ptr, [0, 1, 2])
    slots = _slots(cv["k"])

    def run():
        ops.tc_edge_mlp_aggregate_f32(PQ[:, :pk.Kp], PQ[:, pk.Kp:], pk.Kp, ell, slots, cv["W2hi"],
                                      cv["W2lo"], pk.b, out, coords, [0, 1, 2])
    for _ in range(3):
        run()
    s = [torch.cuda.Event(enable_timing=True) for _ in range(reps)]
    e = [torch.cuda.Event(enable_timing=True) for _ in range(reps)]
    for i in range(reps):
        flush_buf.fill_(i)
        s[i].record()
        run()
        e[i].record()
    torch.cuda.synchronize(dev)
    ms = sum(a.elapsed_time(b_) for a, b_ in zip(s, e)) / reps
    executed = 3 * 2.0 * (N * slots) * pk.Kp * pk.b16     # three MMAs per K step
    name = (f"gn::tc::edge_mlp_aggregate_x3_kernel<{slots},false> (fp32 P / Q / out; three "
            "tcgen05.mma.cta_group::2 kind::f16 per K step on fp16 (hi, lo) operand pairs)")
    return ms, executed, name
