"""Training path of DynEdge on the 16-bit tensor-core kernels (BASELINE config 3).

`dynedge_train_forward(model, data)` runs the same forward as inference
(`dynedge_bf16.forward_bf16`) under a `torch.autograd.Function` whose backward is made of
hand-written kernels (csrc/backward.cu + the tcgen05 GEMM); parameter gradients arrive in
`param.grad` through autograd, so `torch.optim` optimisers and
`torch.nn.parallel.DistributedDataParallel` (NCCL gradient all-reduce) work on the module as they
do on the reference's (`StandardModel.configure_optimizers`, reference
models/easy_model.py:77-96; Lightning DDP strategy, reference easy_model.py:101-124).

What is differentiated: every Linear/ReLU of the EdgeConv MLPs, post-processing and readout, the
"add" aggregation and the global pooling (reference dynedge.py:193-265, 296-350).  The graphs
(kNN indices) carry no gradient, exactly as in the reference, and the model input `x` (detector
pulses, and the global variables derived from it) needs none.
"""

from __future__ import annotations

import os
from typing import Any, Dict, List, Optional, Tuple

import torch
from torch import Tensor

from . import ops
from .dynedge import _parse_mlp
from .dynedge_bf16 import forward_bf16, prepare_training


# GN_TRAIN_GRAPH: 0 direct launches (default), 1 the backward pass as one updated CUDA graph
# (ops.run_captured), 2 the forward too.  Measured on one box (C3, ms per step): 8.4 - 8.6 direct,
# 9.7 - 12.4 level 1, 9.7 - 10.4 level 2: a training step is ~180 launches enqueued from Python in
# about the time the kernels take, and a graph only starts once the whole sequence has been
# captured - the overlap of enqueueing with execution is worth more than the ~2 us per boundary.
_GRAPH_LEVEL = int(os.environ.get("GN_TRAIN_GRAPH", "0"))


def _param_list(model) -> List[torch.nn.Parameter]:
    """Parameters in the order the backward pass returns their gradients."""
    out: List[torch.nn.Parameter] = []
    for conv in model._conv_layers:
        linears, _ = conv._mlp()
        for lin in linears:
            out += [lin.weight, lin.bias]
    for seq in (model._post_processing, model._readout):
        linears, _ = _parse_mlp(seq)
        for lin in linears:
            out += [lin.weight, lin.bias]
    return out


def _t16(x: Tensor) -> Tensor:
    """Small weight matrices: transposed contiguous copy, padded to 8 columns (plumbing)."""
    r, c = x.shape
    out = torch.zeros((c, (r + 7) // 8 * 8), dtype=x.dtype, device=x.device)
    out[:, :r] = x.t()
    return out


def _linear_weight_grad(dZT: Tensor, XT: Tensor, n_out: int, k_in: int, rows: int) -> Tensor:
    """dW [n_out, k_in] = dZ^T X from the transposed operands (K = rows, split K)."""
    dW = ops.zeros((n_out, k_in), torch.float32, dZT.device)
    ops.tc_linear_accumulate(dZT[:n_out], XT[:k_in], dW, K=rows, splits=0)
    return dW


class _DynEdgeTrain(torch.autograd.Function):
    @staticmethod
    def forward(ctx, model, data, *params):  # noqa: D401 - params only tie autograd to the module
        with torch.no_grad():
            prep = prepare_training(model, data)       # tensor ops, weight repacks, one host read

            def body():
                save: Dict[str, Any] = {"layers": [], "post": [], "readout": []}
                return forward_bf16(model, data, save=save, prep=prep), save
            # launches and torch.empty only from here: one updated CUDA graph per step - unless
            # the batch needs tensor ops in between (global variables appended after pooling, a
            # data set graph with more in-edges than slots)
            if _GRAPH_LEVEL >= 2 and prep[3].ell_stride > 0 and \
                    not model._add_global_variables_after_pooling:
                out, save = ops.run_captured(id(model) + 2, prep[0].device, body)
            else:
                out, save = body()
        ctx.model = model
        ctx.saved = save
        return out

    @staticmethod
    def backward(ctx, dout):
        model, sv = ctx.model, ctx.saved
        with torch.no_grad():
            grads = _backward(model, sv, dout.float().contiguous())
        return (None, None) + tuple(grads)


def _backward(model, sv: Dict[str, Any], dout: Tensor) -> List[Optional[Tensor]]:
    """Three parts: (A) the readout, a few [B, .] tensor ops; (B) pooling, post-processing MLP and
    EdgeConv layers - ONLY launches of this library and `torch.empty`, so the whole part replays as
    one updated CUDA graph (`ops.run_captured`: ~120 launches without idle GPU between them);
    (C) slicing / re-ordering of the raw gradient matrices into the parameters' shapes."""
    dev = dout.device
    N, B = sv["N"], sv["B"]
    skip, offs, widths, D0, total = sv["skip"], sv["offs"], sv["widths"], sv["D0"], sv["total"]
    ptr = sv["ptr"]
    g_ro: List[Tensor] = []

    # ------------------------------------------------------------ (A) readout (fp32, [B, .])
    d = dout
    for lin, hin, hout, act in reversed(sv["readout"]):
        if act == "relu":
            d = d * (hout > 0)
        elif act != "none":
            raise NotImplementedError("training supports ReLU readout")
        d = d.contiguous()
        dW = ops.linear_f32(d.t().contiguous(), hin.t().contiguous(), None)      # [n_out, k_in]
        db = d.sum(0)
        d = ops.linear_f32(d, lin.weight.detach().t().contiguous(), None)        # [B, k_in]
        g_ro = [dW, db] + g_ro
    post_outs = sv["post"]
    packs = sv["post_packs"]
    schemes = model._global_pooling_schemes
    if sv["pooled"] is not None:
        d = d[:, :post_outs[-1].shape[1] * len(schemes)].contiguous()
    WTs = [_t16(Wp[:, :K]) for Wp, _b, _n, K in packs]          # [K, n pad 8] (weights only)
    layers = []
    for L in sv["layers"]:
        graph = L["graph"]
        if graph.ell_stride <= 0:   # the data set's edge_index (layer 1): CSR -> slot table
            stride = int(graph.max_degree.item()) if graph.max_degree is not None else 0
            graph = ops.csr_to_ell(graph, max(stride, 1))
        if L["pack"].b % 2:
            raise NotImplementedError("training needs an even EdgeConv output width")
        layers.append((L, graph))
    use_lists = bool(os.environ.get("GN_PQ_LISTS"))

    # ------------------------------------------------------------ (B) launches only
    def body():
        # pooling
        if sv["pooled"] is not None:
            dY = ops.segment_pool_backward(post_outs[-1], ptr, schemes, sv["pooled"], d)
        else:
            dY = d  # node-level readout
        # post-processing MLP
        inputs = [skip] + post_outs[:-1]
        skipT = ops.transpose_16(skip)                   # [total, Np]: X^T of every GEMM on skip
        dskip = None
        raw_post = []
        for i in reversed(range(len(packs))):
            Wp, bias, n, K = packs[i]
            Y = post_outs[i]
            dZ = ops.relu_backward(dY, Y[:, :n])         # bf16 [N, n]
            dZT = ops.transpose_16(dZ)                   # [n, Np]
            XT = skipT if i == 0 else ops.transpose_16(inputs[i][:, :K])
            dW = _linear_weight_grad(dZT, XT, n, K, N)
            db = ops.zeros((n,), torch.float32, dev)
            ops.colsum_accumulate(dZ, db)
            if i == 0:
                dskip = torch.empty((N, total), dtype=torch.float32, device=dev)
                ops.tc_linear_bf16(dZ, WTs[i], None, dskip, K=n, act="none")
            else:
                dX = torch.empty((N, (K + 7) // 8 * 8), dtype=torch.bfloat16, device=dev)
                ops.tc_linear_bf16(dZ, WTs[i], None, dX, K=n, act="none")
                dY = dX[:, :K]
            raw_post.insert(0, (dW, db))
        # EdgeConv layers, last first
        raw_conv = []
        for L, graph in reversed(layers):
            pk = L["pack"]
            P, Q = L["P"], L["Q"]
            Kp, b, b16, d_in = pk.Kp, pk.b, pk.b16, L["d_in"]
            S = graph.ell_stride
            E = N * S
            Ep = (E + 7) // 8 * 8
            mask = L.get("relu_mask")
            dOut = dskip[:, L["off_out"]:L["off_out"] + b]
            # (pad columns feed the GEMMs as zeros)
            alloc = ops.zeros if b16 != b else \
                (lambda shape, dtype, device: torch.empty(shape, dtype=dtype, device=device))
            HT = torch.empty((Kp, Ep), dtype=torch.bfloat16, device=dev)
            G2 = alloc((E, b16), torch.bfloat16, dev)
            G2T = alloc((b16, Ep), torch.bfloat16, dev)
            if mask is not None:
                # the forward recorded the ReLU bits of the second Linear: no Z, no row-major H
                ops.edge_build_h(P, Q, Kp, graph, None, HT)
                ops.edge_grad_z_mask(mask, dOut, b, graph, G2, G2T)
            else:
                H = torch.empty((E, Kp), dtype=torch.bfloat16, device=dev)
                ops.edge_build_h(P, Q, Kp, graph, H, HT)
                Z = torch.empty((E, b16), dtype=torch.bfloat16, device=dev)
                ops.tc_linear_bf16(H, pk.W2b, None, Z, K=Kp, act="none")
                ops.edge_grad_z(Z, dOut, b, graph, G2, G2T)
                del Z, H
            dW2p = ops.zeros((b16, Kp), torch.float32, dev)
            ops.tc_linear_accumulate(G2T, HT, dW2p, K=E, splits=0)
            dH = torch.empty((E, Kp), dtype=torch.bfloat16, device=dev)
            ops.tc_linear_bf16(G2, pk.W2bT, None, dH, K=b16, act="none")
            del G2, G2T, HT
            # dP, dQ: every node sums its own slots (dP, registers) and the slots that name it
            # (dQ: vector reductions into dQf; GN_PQ_LISTS=1: through the graph grouped by source,
            # a fixed summation order at the price of a second read of dH)
            dPQ = torch.empty((N, 2 * Kp), dtype=torch.bfloat16, device=dev)
            dP32 = torch.empty((N, Kp), dtype=torch.float32, device=dev) if pk.first else None
            dQf = torch.empty((N, Kp), dtype=torch.float32, device=dev)   # fp32 accumulator of dQ
            rev = ops.ell_reverse(graph, ptr) if use_lists else None
            ops.edge_grad_pq_gather(dH, P, Q, Kp, graph, ptr, rev, dPQ[:, :Kp], dPQ[:, Kp:], dP32,
                                    dQf)
            del dH
            db1 = ops.zeros((Kp,), torch.float32, dev)
            ops.colsum_accumulate(dPQ[:, :Kp], db1)
            off_in = L["off_in"]
            dWpq = ops.zeros((2 * Kp, d_in), torch.float32, dev)
            if pk.first:   # raw input features: fp32 throughout (see gn_outer_accumulate_f32)
                ops.outer_accumulate_f32(dP32, sv["x0"], dWpq[:Kp])
                ops.outer_accumulate_f32(dQf, sv["x0"], dWpq[Kp:])
            else:
                dPQT = ops.transpose_16(dPQ)                                # [2Kp, Np]
                ops.tc_linear_accumulate(dPQT, skipT[off_in:off_in + d_in], dWpq, K=N, splits=0)
                # gradient with respect to the layer input: onto its skip slice
                ops.tc_linear_accumulate(dPQ, pk.WpqT[:d_in], dskip[:, off_in:], K=2 * Kp,
                                         splits=1)
            raw_conv.insert(0, (dWpq, db1, dW2p))
        return raw_post, raw_conv

    if use_lists or _GRAPH_LEVEL < 1:   # (ell_reverse is built with tensor ops)
        raw_post, raw_conv = body()
    else:
        raw_post, raw_conv = ops.run_captured(id(model) + 1, dev, body)

    # ------------------------------------------------------------ (C) parameter shapes
    g_post: List[Optional[Tensor]] = []
    for i, (dW, db) in enumerate(raw_post):
        if i == 0:  # un-scatter the padded skip layout back to the reference's column order
            # (a device index built once: `dW[:, python_list]` converts the 1043-entry list on
            # the host at every call - 1.2 ms of a step that follows the host)
            key = ("_unscatter_idx", str(dev), D0, tuple(widths), tuple(offs))
            idx = model.__dict__.get("_unscatter_cache", (None, None))
            if idx[0] != key:
                cols = list(range(D0))
                for w, o in zip(widths, offs):
                    cols += list(range(o, o + w))
                idx = (key, torch.tensor(cols, dtype=torch.long, device=dev))
                model.__dict__["_unscatter_cache"] = idx
            dW = dW.index_select(1, idx[1])
        g_post += [dW, db if packs[i][1] is not None else None]
    flat: List[Optional[Tensor]] = []
    for li, (dWpq, db1, dW2p) in enumerate(raw_conv):
        pk = sv["layers"][li]["pack"]
        Kp, a, b = pk.Kp, pk.a, pk.b
        dWp, dWq = dWpq[:a], dWpq[Kp:Kp + a]
        dW1 = torch.cat([dWp, dWq - dWp], 1)                                # [a, 2d]
        l1, l2 = model._conv_layers[li]._mlp()[0]
        flat += [dW1, db1[:a] if l1.bias is not None else None,
                 dW2p[:b, :a].contiguous(),
                 dW2p[:b, a].contiguous() if l2.bias is not None else None]
    return flat + g_post + g_ro


def dynedge_train_forward(model, data: Any) -> Tensor:
    """DynEdge forward that records what its hand-written backward needs (autograd enabled)."""
    if model.precision not in ("bf16",):
        raise NotImplementedError("training runs on the 16-bit path: DynEdge(precision='bf16')")
    params = _param_list(model)
    return _DynEdgeTrain.apply(model, data, *params)
