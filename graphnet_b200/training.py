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
from .dynedge_bf16 import forward_bf16


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
    dW = torch.zeros((n_out, k_in), dtype=torch.float32, device=dZT.device)
    ops.tc_linear_accumulate(dZT[:n_out], XT[:k_in], dW, K=rows, splits=0)
    return dW


class _DynEdgeTrain(torch.autograd.Function):
    @staticmethod
    def forward(ctx, model, data, *params):  # noqa: D401 - params only tie autograd to the module
        save: Dict[str, Any] = {"layers": [], "post": [], "readout": []}
        with torch.no_grad():
            out = forward_bf16(model, data, save=save)
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
    dev = dout.device
    N, B = sv["N"], sv["B"]
    skip, offs, widths, D0, total = sv["skip"], sv["offs"], sv["widths"], sv["D0"], sv["total"]
    ptr = sv["ptr"]
    Np = (N + 7) // 8 * 8
    g_conv: List[List[Tensor]] = []
    g_post: List[Tensor] = []
    g_ro: List[Tensor] = []

    # ---------------------------------------------------------------- readout (fp32, [B, .])
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

    # ---------------------------------------------------------------- pooling
    post_outs = sv["post"]
    Y2 = post_outs[-1]
    if sv["pooled"] is not None:
        schemes = model._global_pooling_schemes
        C_ = Y2.shape[1]
        dY = ops.segment_pool_backward(Y2, ptr, schemes, sv["pooled"], d[:, :C_ * len(schemes)]
                                       .contiguous())
    else:
        dY = d  # node-level readout

    # ---------------------------------------------------------------- post-processing MLP
    packs = sv["post_packs"]
    inputs = [skip] + post_outs[:-1]
    skipT = ops.transpose_16(skip)                       # [total, Np]: X^T of every GEMM on skip
    dskip = None
    for i in reversed(range(len(packs))):
        Wp, bias, n, K = packs[i]
        Y = post_outs[i]
        dZ = ops.relu_backward(dY, Y[:, :n])             # bf16 [N, n]
        dZT = ops.transpose_16(dZ)                       # [n, Np]
        XT = skipT if i == 0 else ops.transpose_16(inputs[i][:, :K])
        dW = _linear_weight_grad(dZT, XT, n, K, N)
        db = torch.zeros(n, dtype=torch.float32, device=dev)
        ops.colsum_accumulate(dZ, db)
        WT = _t16(Wp[:, :K])                             # [K, n pad 8]
        if i == 0:
            dskip = torch.empty((N, total), dtype=torch.float32, device=dev)
            ops.tc_linear_bf16(dZ, WT, None, dskip, K=n, act="none")
        else:
            dX = torch.empty((N, (K + 7) // 8 * 8), dtype=torch.bfloat16, device=dev)
            ops.tc_linear_bf16(dZ, WT, None, dX, K=n, act="none")
            dY = dX[:, :K]
        if i == 0:  # un-scatter the padded skip layout back to the reference's column order
            cols = list(range(D0))
            for w, o in zip(widths, offs):
                cols += list(range(o, o + w))
            dW = dW[:, cols]
        g_post = [dW, db if bias is not None else None] + g_post

    # ---------------------------------------------------------------- EdgeConv layers, last first
    for li in reversed(range(len(sv["layers"]))):
        L = sv["layers"][li]
        pk, graph = L["pack"], L["graph"]
        P, Q = L["P"], L["Q"]
        Kp, a, b, b16, d_in = pk.Kp, pk.a, pk.b, pk.b16, L["d_in"]
        if graph.ell_stride <= 0:   # the data set's edge_index (layer 1): CSR -> slot table
            stride = int(graph.max_degree.item()) if graph.max_degree is not None else 0
            graph = ops.csr_to_ell(graph, max(stride, 1))
        S = graph.ell_stride
        E = N * S
        Ep = (E + 7) // 8 * 8
        mask = L.get("relu_mask")
        if b % 2:
            raise NotImplementedError("training needs an even EdgeConv output width")
        dOut = dskip[:, L["off_out"]:L["off_out"] + b]
        alloc = torch.zeros if b16 != b else torch.empty   # pad columns feed the GEMMs as zeros
        HT = torch.empty((Kp, Ep), dtype=torch.bfloat16, device=dev)
        W2b = pk.W2b                                                        # [b16, Kp] bf16
        G2 = alloc((E, b16), dtype=torch.bfloat16, device=dev)
        G2T = alloc((b16, Ep), dtype=torch.bfloat16, device=dev)
        if mask is not None:
            # the forward recorded the ReLU bits of the second Linear: no Z, no row-major H
            ops.edge_build_h(P, Q, Kp, graph, None, HT)
            ops.edge_grad_z_mask(mask, dOut, b, graph, G2, G2T)
        else:
            H = torch.empty((E, Kp), dtype=torch.bfloat16, device=dev)
            ops.edge_build_h(P, Q, Kp, graph, H, HT)
            Z = torch.empty((E, b16), dtype=torch.bfloat16, device=dev)
            ops.tc_linear_bf16(H, W2b, None, Z, K=Kp, act="none")
            ops.edge_grad_z(Z, dOut, b, graph, G2, G2T)
            del Z, H
        dW2p = torch.zeros((b16, Kp), dtype=torch.float32, device=dev)
        ops.tc_linear_accumulate(G2T, HT, dW2p, K=E, splits=0)
        dH = torch.empty((E, Kp), dtype=torch.bfloat16, device=dev)
        ops.tc_linear_bf16(G2, pk.W2bT, None, dH, K=b16, act="none")
        del G2, G2T, HT
        # dP, dQ: every node sums its own slots (dP, registers) and the slots that name it (dQ:
        # vector reductions into dQf; GN_PQ_LISTS=1: through the graph grouped by source, a
        # fixed summation order at the price of a second read of dH)
        dPQ = torch.empty((N, 2 * Kp), dtype=torch.bfloat16, device=dev)
        dP32 = torch.empty((N, Kp), dtype=torch.float32, device=dev) if pk.first else None
        dQf = torch.empty((N, Kp), dtype=torch.float32, device=dev)   # fp32 accumulator of dQ
        rev = ops.ell_reverse(graph, ptr) if os.environ.get("GN_PQ_LISTS") else None
        ops.edge_grad_pq_gather(dH, P, Q, Kp, graph, ptr, rev, dPQ[:, :Kp], dPQ[:, Kp:], dP32, dQf)
        del dH
        db1 = torch.zeros(Kp, dtype=torch.float32, device=dev)
        ops.colsum_accumulate(dPQ[:, :Kp], db1)
        off_in = L["off_in"]
        dWpq = torch.zeros((2 * Kp, d_in), dtype=torch.float32, device=dev)
        if pk.first:   # raw input features: fp32 throughout (see gn_outer_accumulate_f32)
            ops.outer_accumulate_f32(dP32, sv["x0"], dWpq[:Kp])
            ops.outer_accumulate_f32(dQf, sv["x0"], dWpq[Kp:])
        else:
            dPQT = ops.transpose_16(dPQ)                                    # [2Kp, Np]
            ops.tc_linear_accumulate(dPQT, skipT[off_in:off_in + d_in], dWpq, K=N, splits=0)
        if not pk.first:  # gradient with respect to the layer input: onto its skip slice
            ops.tc_linear_accumulate(dPQ, pk.WpqT[:d_in], dskip[:, off_in:], K=2 * Kp, splits=1)
        dWp, dWq = dWpq[:a], dWpq[Kp:Kp + a]
        dW1 = torch.cat([dWp, dWq - dWp], 1)                                # [a, 2d]
        linears, _ = model._conv_layers[li]._mlp()
        l1, l2 = linears
        g = [dW1, db1[:a] if l1.bias is not None else None,
             dW2p[:b, :a].contiguous(), dW2p[:b, a].contiguous() if l2.bias is not None else None]
        g_conv.insert(0, g)

    flat: List[Optional[Tensor]] = []
    for g in g_conv:
        flat += g
    return flat + g_post + g_ro


def dynedge_train_forward(model, data: Any) -> Tensor:
    """DynEdge forward that records what its hand-written backward needs (autograd enabled)."""
    if model.precision not in ("bf16",):
        raise NotImplementedError("training runs on the 16-bit path: DynEdge(precision='bf16')")
    params = _param_list(model)
    return _DynEdgeTrain.apply(model, data, *params)
