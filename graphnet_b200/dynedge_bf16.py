"""16-bit tensor-core forward of DynEdge (tcgen05 / TMEM / TMA kernels).

Same computation as `DynEdge._forward_f32` (reference dynedge.py:296-350) with
16-bit GEMM operands and fp32 accumulation / aggregation; outputs agree with the
fp32 reference within 2e-2 relative (BASELINE.json north_star).

precision="bf16": layer inputs/outputs (the skip-cat buffer) and the node-GEMM /
post-processing weights are bf16; precision="fp16": the same tensors are IEEE
fp16 (3 more mantissa bits, conversions saturate at +-65504).  In both modes the
raw input features enter the first EdgeConv in fp32 (K = 2F+5 is tiny) and the
hidden layer of every EdgeConv MLP (P, Q, h, W2) is fp16 - see csrc/tc_edge.cu.
Weight repacking (padding, conversion, bias folding) happens once per parameter
version on the device with torch ops - plumbing, not the timed path.
"""

from __future__ import annotations

import ctypes as C
from typing import Any, Dict, List, Optional, Tuple

import torch
from torch import Tensor

from . import ops
from .dynedge import _act_name, _parse_mlp, _WeightCache


def _round_up(v: int, m: int) -> int:
    return (v + m - 1) // m * m


def _slots(k: int) -> int:
    """Slots per node of the edge kernel: the smallest supported count >= k+1 (in-degree of a
    kNN graph is at most k+1), so no second pass is needed."""
    for s in (8, 9, 16, 17, 32):
        if s >= k + 1:
            return s
    return 32


def _slots_min(k: int) -> int:
    """`DynEdge.min_edge_slots = True` (inference, aggr='add'): the smallest supported slot count
    >= k.  A kNN row holds k + 1 neighbours only where more than k points coincide with the query;
    those entries go through the kernel's overflow pass and every tile row carries an edge (k = 16:
    4 nodes per 64-row tile instead of 3, k = 8: 8 instead of 7).  Measured at C2 / C4: the main
    pass gets 7 % / 25 % shorter, but a layer whose kNN columns are tied for most nodes (ReLU-dead
    coordinate columns of a randomly initialised model: 59 % of the nodes of one C2 layer) pays
    4 x for its overflow pass - off by default."""
    for s in (8, 9, 16, 17, 32):
        if s >= k:
            return s
    return 32


class _ConvPack:
    """Repacked weights of one DynEdgeConv MLP for the tensor-core kernels."""

    def __init__(self, l1: torch.nn.Linear, l2: torch.nn.Linear, io_dtype: torch.dtype,
                 first: bool, train: bool = False, act: str = "relu"):
        a, d2 = l1.weight.shape
        d = d2 // 2
        b = l2.weight.shape[0]
        Kp = _round_up(a + 2, 32)
        if Kp > 352 or b > 256:
            raise NotImplementedError(
                f"bf16 EdgeConv kernel supports hidden <= 350 and output <= 256 (got {a}, {b}); "
                "use precision='fp32' for this architecture")
        Kd = d if first else _round_up(d, 8)
        self.d, self.a, self.b, self.Kp, self.Kd, self.b16 = d, a, b, Kp, Kd, _round_up(b, 16)
        # first layer: fp32 FFMA node GEMM on the fp32 input features; later layers: tcgen05
        self.first = first
        # one kernel writes every layout (csrc/backward.cu: edgeconv_pack_tc_kernel): Wpq = [P | Q]
        # weights (P = x Wp^T + bias_p, Q = x Wq^T), the two "ones" columns that carry b2 through
        # the second GEMM, W2p in fp16 (the hidden layer runs in fp16, see csrc/tc_edge.cu) and,
        # in training, the bf16 / transposed copies the backward GEMMs read
        self.act = act
        packed = ops.edgeconv_pack_tc(l1.weight, l1.bias, l2.weight, l2.bias, Kp, Kd,
                                      torch.float32 if first else io_dtype, train, act=act)
        self.Wpq, self.bias_pq, self.W2p, self.W2b, self.W2bT, self.WpqT = packed
        self.Wp, self.Wq = self.Wpq[:Kp], self.Wpq[Kp:]
        self.bias_p = self.bias_pq[:Kp]
        self._split = None

    def split_weights(self, kc: int) -> Tensor:
        """First layer: Wpq (fp32 [2Kp, d]) as bf16 [2Kp, 3*kc] = [W_hi | W_hi | W_lo], the partner
        of a GN_BF16X3 input ([x_hi | x_lo | x_hi]); built once per weight version (plumbing)."""
        if self._split is None or self._split.shape[1] != 3 * kc:
            W = self.Wpq.float()
            hi = W.to(torch.bfloat16)
            lo = (W - hi.float()).to(torch.bfloat16)
            out = torch.zeros((W.shape[0], 3 * kc), dtype=torch.bfloat16, device=W.device)
            d = W.shape[1]
            out[:, :d] = hi
            out[:, kc:kc + d] = hi
            out[:, 2 * kc:2 * kc + d] = lo
            self._split = out.contiguous()
        return self._split


def _io_dtype(model) -> torch.dtype:
    return torch.float16 if model.precision == "fp16" else torch.bfloat16


def _conv_pack(conv, io_dtype: torch.dtype, first: bool, train: bool = False) -> _ConvPack:
    linears, act = conv._mlp()
    if act not in ("relu", "gelu", "leaky_relu") or len(linears) != 2:
        raise NotImplementedError(
            "the 16-bit path supports Linear-act-Linear-act EdgeConv MLPs with ReLU, GELU or "
            "LeakyReLU; use precision='fp32'")
    if conv.aggr not in ("add", "sum", "mean", "max", "min"):
        raise NotImplementedError(f"aggr={conv.aggr!r}: add, mean, max or min")
    if train and (act != "relu" or conv.aggr not in ("add", "sum")):
        raise NotImplementedError("training runs on ReLU / aggr='add' EdgeConv layers")
    key = f"_cache_tc_{io_dtype}_{first}_{train}_{act}"
    if not hasattr(conv, key):
        setattr(conv, key, _WeightCache())
    l1, l2 = linears
    params = [p for p in (l1.weight, l1.bias, l2.weight, l2.bias) if p is not None]
    return getattr(conv, key).get(params, lambda: _ConvPack(l1, l2, io_dtype, first, train, act))


def _skip_layout(model) -> Tuple[int, List[int], List[int], int]:
    """Column layout of the bf16 skip-cat buffer: [x0 | pad | conv1 | pad | ...]."""
    D0 = model._nb_conv_inputs
    widths = [conv.out_features for conv in model._conv_layers]
    offs, off = [], _round_up(D0, 8)
    for w in widths:
        offs.append(off)
        off += _round_up(w, 8)
    return D0, widths, offs, off


def _post_pack(model, total: int, D0: int, widths: List[int], offs: List[int],
               io_dtype: torch.dtype):
    key = f"_cache_post_{io_dtype}"
    if not hasattr(model, key):
        setattr(model, key, _WeightCache())
    linears, act = _parse_mlp(model._post_processing)
    if act not in ("relu", "gelu"):
        raise NotImplementedError("the 16-bit path supports ReLU and GELU; use precision='fp32'")
    params = [p for lin in linears for p in (lin.weight, lin.bias) if p is not None]

    def build():
        packs = []
        dev = linears[0].weight.device
        for i, lin in enumerate(linears):
            W = lin.weight.detach().float()
            n, k = W.shape
            if i == 0:  # scatter the columns to the padded skip layout
                Wp = torch.zeros((n, total), dtype=torch.float32, device=dev)
                Wp[:, :D0] = W[:, :D0]
                src = D0
                for w, o in zip(widths, offs):
                    Wp[:, o:o + w] = W[:, src:src + w]
                    src += w
                K = total
            else:
                K = k
                Wp = torch.zeros((n, _round_up(k, 8)), dtype=torch.float32, device=dev)
                Wp[:, :k] = W
            bias = None if lin.bias is None else lin.bias.detach().float().contiguous()
            packs.append((Wp.to(io_dtype).contiguous(), bias, n, K))
        return packs
    return getattr(model, key).get(params, build)


class _NativePlan:
    """`gn_dynedge_plan` of a model (csrc/forward.cu) plus the tensors its pointers refer to."""

    def __init__(self, model, io: torch.dtype):
        from . import _lib
        D0, widths, offs, total = _skip_layout(model)
        convs = list(model._conv_layers)
        posts = _post_pack(model, total, D0, widths, offs, io)
        pact = _parse_mlp(model._post_processing)[1]
        ro, ract = _parse_mlp(model._readout)
        pools = model._global_pooling_schemes
        if len(convs) > _lib.GN_MAX_LAYERS or len(posts) > _lib.GN_MAX_DENSE or len(ro) != 1 \
                or pact != "relu" or not pools or posts[-1][2] % 2:
            raise NotImplementedError("outside the native driver's shapes")
        plan = _lib.GnDynedgePlan()
        keep: List[Any] = []
        x0w = offs[0]
        plan.F, plan.x0w, plan.skip_cols = model._nb_inputs, x0w, total
        plan.io_dtype = ops._DT[io]
        plan.n_conv, plan.n_post, plan.n_readout, plan.n_pools = len(convs), len(posts), 1, len(pools)
        plan.input_k = model.input_graph_neighbours or model._nb_neighbours
        off_in, d_in = 0, D0
        for li, conv in enumerate(convs):
            if conv.aggr not in ("add", "sum"):
                raise NotImplementedError("native driver: aggr='add'")
            pk = _conv_pack(conv, io, first=(li == 0))
            assert pk.d == d_in
            W = pk.split_weights(x0w) if li == 0 else pk.Wpq
            cols = sorted(ops._columns(conv.features_subset, pk.b))
            if len(cols) > 16:
                raise NotImplementedError("native driver: at most 16 kNN columns")
            c = plan.conv[li]
            c.Wpq, c.bias_pq, c.W2p = W.data_ptr(), pk.bias_pq.data_ptr(), pk.W2p.data_ptr()
            c.d_in, c.Kp, c.b, c.k = d_in, pk.Kp, pk.b, conv.nb_neighbors
            c.slots = (_slots_min if model.min_edge_slots else _slots)(conv.nb_neighbors)
            c.act = ops.ACT[pk.act]
            c.off_in, c.off_out, c.ncoord, c.ldw = off_in, offs[li], len(cols), W.stride(0)
            for t, col in enumerate(cols):
                c.coord_cols[t] = col
            keep += [pk, W]
            off_in, d_in = offs[li], pk.b
        for i, (Wp, bias, n, K) in enumerate(posts):
            d = plan.post[i]
            d.W, d.bias = Wp.data_ptr(), 0 if bias is None else bias.data_ptr()
            d.n, d.K, d.act, d.ldw, d.tc = n, K, ops.ACT[pact], Wp.stride(0), 1
            keep += [Wp, bias]
        lin = ro[0]
        K_ = posts[-1][2] * len(pools)
        if lin.in_features != K_:
            raise NotImplementedError("native driver: readout width")
        tc = K_ % 8 == 0 and K_ >= 256 and ract in ("relu", "none")
        if tc:  # fp16 operands, weights times 2^6 and the pooled input times 2^-6 (exact)
            Wr = (lin.weight.detach().float() * 64.0).to(torch.float16).contiguous()
        else:
            Wr = lin.weight.detach().float().contiguous()
        br = None if lin.bias is None else lin.bias.detach().float().contiguous()
        d = plan.readout[0]
        d.W, d.bias = Wr.data_ptr(), 0 if br is None else br.data_ptr()
        d.n, d.K, d.act, d.ldw, d.tc = lin.out_features, K_, ops.ACT[ract], Wr.stride(0), int(tc)
        keep += [Wr, br]
        for i, s in enumerate(pools):
            plan.pools[i] = ops.POOL[s]
        self.plan, self.keep, self.n_out = plan, keep, lin.out_features
        self.ws_bytes: Dict[Tuple[int, int, int, bool, int], int] = {}


def _native_plan(model, io: torch.dtype):
    """The cached plan, or None when the model is outside what csrc/forward.cu drives (the Python
    sequence below then runs the same kernels one call at a time)."""
    key = f"_cache_native_{io}_{int(model.min_edge_slots)}"
    cache = model.__dict__.get(key)
    if cache is None:
        cache = _WeightCache()
        model.__dict__[key] = cache
    if model._add_global_variables_after_pooling or model._skip_readout:
        return None
    params = list(model.parameters())

    def build():
        try:
            return _NativePlan(model, io)
        except NotImplementedError:
            return None
    return cache.get(params, build)


def _forward_native(model, np_: _NativePlan, data: Any, edge_index: Optional[Tensor]) -> Tensor:
    """One C-ABI call (gn_dynedge_infer16) enqueues the whole forward over one workspace."""
    from . import _lib
    x = data.x
    if not x.is_cuda:
        raise RuntimeError("graphnet_b200.DynEdge needs CUDA tensors (no CPU fallback)")
    dev = x.device
    if x.dtype != torch.float32 or x.stride(1) != 1:
        x = x.float().contiguous()
    N = x.shape[0]
    assert x.shape[1] == model._nb_inputs, \
        f"expected {model._nb_inputs} input features, got {x.shape[1]}"
    n_pulses = data.n_pulses.reshape(-1)
    B = n_pulses.shape[0]
    if n_pulses.dtype != torch.int32 or not n_pulses.is_contiguous():
        n_pulses = n_pulses.to(torch.int32).contiguous()
    batch = data.batch
    if not batch.is_contiguous():
        batch = batch.contiguous()
    assert batch.shape[0] == N
    if edge_index is None:
        edge_index = getattr(data, "edge_index", None)
    E = 0
    if edge_index is not None:
        assert edge_index.dim() == 2 and edge_index.shape[0] == 2
        if edge_index.dtype != torch.int64 or not edge_index.is_contiguous():
            edge_index = edge_index.long().contiguous()
        E = edge_index.shape[1]
    lib = _lib.load()
    # (not a weight: may be changed between calls)
    input_k = model.input_graph_neighbours or model._nb_neighbours
    np_.plan.input_k = input_k
    plan = C.byref(np_.plan)
    wkey = (N, B, E, edge_index is not None, input_k)
    nbytes = np_.ws_bytes.get(wkey)
    if nbytes is None:
        nbytes = int(lib.gn_dynedge_infer16_workspace(plan, N, B, E, int(edge_index is not None)))
        if len(np_.ws_bytes) > 4096:
            np_.ws_bytes.clear()
        np_.ws_bytes[wkey] = nbytes
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    out = torch.empty((B, np_.n_out), dtype=torch.float32, device=dev)
    flags = torch.empty(1, dtype=torch.int32, device=dev)
    with ops._on(dev):
        ops.check(lib.gn_dynedge_infer16(
            plan, x.data_ptr(), x.stride(0), batch.data_ptr(), n_pulses.data_ptr(), N, B,
            None if edge_index is None else edge_index.data_ptr(), E, ws.data_ptr(), nbytes,
            out.data_ptr(), out.stride(0), flags.data_ptr(), ops._stream(dev)),
            "gn_dynedge_infer16")
    model._last_flags = flags
    return out


def _readout_weight(model, ri: int, lin) -> Tensor:
    """fp16 copy of a readout Linear's weight times 2^6 (per parameter version)."""
    key = f"_cache_readout_{ri}"
    if not hasattr(model, key):
        setattr(model, key, _WeightCache())
    return getattr(model, key).get(
        [lin.weight],
        lambda: (lin.weight.detach().float() * 64.0).to(torch.float16).contiguous())


def prepare_training(model, data: Any):
    """Everything of the training forward that is NOT a launch of this library: the batch's graph
    inputs (tensor fills inside), the data set's CSR graph as a slot table (one device -> host
    read), the weight repacks of this parameter version (tensor ops on a cache miss, i.e. every
    step).  `forward_bf16(..., save=..., prep=...)` is launches and `torch.empty` only after this
    and can replay as one updated CUDA graph (`ops.run_captured`)."""
    x, n_pulses, ptr, graph, N, B = model._graph_inputs(data, None)
    io = _io_dtype(model)
    convs = list(model._conv_layers)
    for li, conv in enumerate(convs):
        _conv_pack(conv, io, first=(li == 0), train=True)
    D0, widths, offs, total = _skip_layout(model)
    _post_pack(model, total, D0, widths, offs, io)
    if not model._skip_readout:
        for ri, lin in enumerate(_parse_mlp(model._readout)[0]):
            _readout_weight(model, ri, lin)
    slots = _slots(convs[0].nb_neighbors)
    if graph.ell_stride <= 0 and graph.max_degree is not None and \
            int(graph.max_degree.item()) <= slots:
        flags = graph.flags
        graph = ops.csr_to_ell(graph, slots)
        graph.flags = flags
    n_pulses = n_pulses.to(torch.int32).contiguous()
    return x, n_pulses, ptr, graph, N, B


def forward_bf16(model, data: Any, edge_index: Optional[Tensor] = None,
                 return_intermediates: bool = False, save: Optional[Dict[str, Any]] = None,
                 prep: Optional[tuple] = None):
    """`save` (training): receives the tensors the backward pass needs (training.py)."""
    if save is None and not return_intermediates and getattr(model, "native_driver", True) and \
            isinstance(getattr(data, "batch", None), Tensor) and data.batch.dtype == torch.int64 \
            and data.x.shape[0] > 0 and data.n_pulses.numel() > 0:
        np_ = _native_plan(model, _io_dtype(model))
        if np_ is not None:
            return _forward_native(model, np_, data, edge_index)
    x, n_pulses, ptr, graph, N, B = prep if prep is not None else \
        model._graph_inputs(data, edge_index)
    model._last_flags = graph.flags
    dev = x.device
    F = model._nb_inputs
    assert x.shape[1] == F, f"expected {F} input features, got {x.shape[1]}"
    io = _io_dtype(model)
    D0, widths, offs, total = _skip_layout(model)
    needs_zero = any(w % 8 for w in widths)
    skip = ops.zeros((N, total), io, dev) if needs_zero else \
        torch.empty((N, total), dtype=io, device=dev)
    x0w = offs[0]
    # inference: the first node GEMM runs on the tensor cores from a bf16 [hi | lo | hi] split of
    # the fp32 input features (GN_BF16X3: 16 mantissa bits, no 8-bit rounding of the coordinates);
    # training keeps the fp32 copy its weight-gradient kernel reads
    x0_split = None
    if model._add_global_variables_after_pooling:
        gv = ops.global_variables(x, graph, ptr, n_pulses)
        skip[:, :x0w] = 0
        skip[:, :F] = x
        x0 = x
    elif save is None:
        x0 = None
        x0_split = torch.empty((N, 3 * x0w), dtype=torch.bfloat16, device=dev)
        gv = ops.global_variables(x, graph, ptr, n_pulses, x0=x0_split, x0_cols=x0w,
                                  x0b=skip, x0b_cols=x0w, x0_split=True)
    else:
        x0 = torch.empty((N, D0), dtype=torch.float32, device=dev)
        gv = ops.global_variables(x, graph, ptr, n_pulses, x0=x0, x0_cols=D0,
                                  x0b=skip, x0b_cols=x0w)
    inter: Dict[str, Any] = {"global_variables": gv, "graphs": []}

    off_in, d_in = 0, D0
    convs = list(model._conv_layers)
    for li, conv in enumerate(convs):
        pk = _conv_pack(conv, io, first=(li == 0), train=save is not None)
        assert pk.d == d_in
        # one node GEMM writes [P | Q] (the layer input is read once); the edge kernel takes the
        # two halves as strided matrices
        PQ = torch.empty((N, 2 * pk.Kp), dtype=torch.float16, device=dev)
        P, Q = PQ[:, :pk.Kp], PQ[:, pk.Kp:]
        aggr = "add" if conv.aggr == "sum" else conv.aggr
        if aggr != "add" and graph.ell_stride <= 0:
            # mean / max / min take every in-edge of a node from one row of slots: the data set's
            # edge_index (layer 1, CSR) goes to the slot-table layout if its degrees allow
            if graph.max_degree is None or int(graph.max_degree.item()) > _slots(conv.nb_neighbors):
                raise NotImplementedError(
                    f"aggr={aggr!r} on the 16-bit path needs in-degrees <= "
                    f"{_slots(conv.nb_neighbors)} (a kNN graph of the layer's k); use precision='fp32'")
            graph = ops.csr_to_ell(graph, _slots(conv.nb_neighbors))
        if li == 0 and x0_split is not None:
            ops.tc_linear_bf16(x0_split, pk.split_weights(x0w), pk.bias_pq, PQ, K=3 * x0w,
                               act="none")
        elif li == 0:
            ops.linear_f32(x0, pk.Wpq, pk.bias_pq, act="none", out=PQ)
        else:
            ops.tc_linear_bf16(skip[:, off_in:], pk.Wpq, pk.bias_pq, PQ, K=d_in, act="none")
        slots = _slots_min(conv.nb_neighbors) if (model.min_edge_slots and save is None
                                                  and aggr == "add") else _slots(conv.nb_neighbors)
        relu_mask = None
        if save is not None:
            if graph.ell_stride <= 0 and graph.max_degree is not None and \
                    int(graph.max_degree.item()) <= slots:
                # the data set's edge_index (layer 1) as a slot table, the layout the backward
                # kernels address edges in
                graph = ops.csr_to_ell(graph, slots)
            if graph.ell_stride in (8, 9, 16, 17, 32):
                relu_mask = ops.edge_relu_mask(N, graph.ell_stride, dev)
            save["layers"].append({"P": P, "Q": Q, "graph": graph, "pack": pk, "off_in": off_in,
                                   "d_in": d_in, "off_out": offs[li], "relu_mask": relu_mask})
        last = li + 1 == len(convs)
        need_graph = (not last) or return_intermediates
        cols = sorted(ops._columns(conv.features_subset, pk.b))
        coords = (torch.empty((N, len(cols)), dtype=torch.float32, device=dev)
                  if need_graph else None)
        ops.tc_edge_mlp_aggregate(P, Q, pk.Kp, graph, slots, pk.W2p, pk.b,
                                  skip[:, offs[li]:], coords, cols, relu_mask=relu_mask,
                                  act=pk.act, aggr=aggr)
        if need_graph:
            graph = ops.knn_graph_ell(coords, conv.nb_neighbors, ptr, list(range(len(cols))))
            if return_intermediates:
                inter["graphs"].append(graph)
                inter.setdefault("coords", []).append(coords)
        off_in, d_in = offs[li], pk.b

    packs = _post_pack(model, total, D0, widths, offs, io)
    pact = _parse_mlp(model._post_processing)[1]
    if save is not None and pact != "relu":
        raise NotImplementedError("training runs on ReLU post-processing layers")
    h = skip
    pools = model._global_pooling_schemes if not model._skip_readout else None
    # inference: the last post-processing Linear + ReLU is fused with the global pooling (its
    # [N, 256] output and the pooling kernel that re-read it never exist); training and
    # return_intermediates keep the separate kernels (the backward pass needs that matrix)
    fuse_pool = bool(pools) and pact == "relu" and save is None and not return_intermediates and \
        isinstance(getattr(data, "batch", None), Tensor) and data.batch.dtype == torch.int64 \
        and packs[-1][2] % 2 == 0
    pooled = hb16 = None
    for i, (Wp, bias, n, K) in enumerate(packs):
        lastp = i + 1 == len(packs)
        if lastp and fuse_pool:
            P = len(pools)
            extra = model._nb_global_variables if model._add_global_variables_after_pooling else 0
            pooled = torch.empty((B, n * P + extra), dtype=torch.float32, device=dev)
            want16 = (n * P + extra) % 8 == 0 and (n * P + extra) >= 256 and not extra
            hb16 = (torch.empty((B, n * P), dtype=torch.float16, device=dev) if want16 else None)
            ops.tc_linear_pool(h, Wp, bias, K, data.batch.contiguous(), ptr, pools, pooled,
                               out16=hb16, scale16=1.0 / 64.0)
            if extra:
                pooled[:, n * P:] = gv
            h = pooled
            break
        out = torch.empty((N, n if lastp else _round_up(n, 8)),
                          dtype=torch.float32 if lastp else io, device=dev)
        ops.tc_linear_bf16(h, Wp, bias, out, K=K, act=pact)
        h = out
        if save is not None:
            save["post"].append(out)
    if pooled is None:
        inter["post"] = h

    if not model._skip_readout:
        if pools and pooled is None:
            P = len(pools)
            C_ = h.shape[1]
            extra = model._nb_global_variables if model._add_global_variables_after_pooling else 0
            pooled = torch.empty((B, C_ * P + extra), dtype=torch.float32, device=dev)
            ops.segment_pool(h, ptr, pools, out=pooled)
            if extra:
                pooled[:, C_ * P:] = gv
            h = pooled
        if pooled is not None:
            inter["pooled"] = pooled
        ro, ract = _parse_mlp(model._readout)
        for ri, lin in enumerate(ro):
            hin = h
            M_, K_ = h.shape
            if K_ % 8 == 0 and K_ >= 256 and ract in ("relu", "none"):
                # the [B, 1024] x [1024, 128] readout on the tensor cores (bf16 operands,
                # fp32 accumulate and output) - the FFMA kernel needs 35 us for it, this 10
                # fp16 operands, the input scaled by 2^-6 and the weights by 2^6 (exact): 11
                # mantissa bits and head-room for sum-pooled values up to 4e6
                Wb = _readout_weight(model, ri, lin)
                hb = hb16 if (ri == 0 and hb16 is not None) else \
                    ops.convert_scale_16(h, torch.float16, 1.0 / 64.0)
                out = torch.empty((M_, lin.out_features), dtype=torch.float32, device=dev)
                bias = None if lin.bias is None else lin.bias.detach().float().contiguous()
                h = ops.tc_linear_bf16(hb, Wb, bias, out, K=K_, act=ract)
            else:
                h = ops.linear_f32(h, lin.weight, lin.bias, act=ract)
            if save is not None:
                save["readout"].append((lin, hin, h, ract))
    if save is not None:
        save.update({"x0": x0, "skip": skip, "offs": offs, "widths": widths, "D0": D0, "total": total,
                     "ptr": ptr, "N": N, "B": B, "post_packs": packs,
                     "pooled": inter.get("pooled")})
    if return_intermediates:
        inter["skip"] = skip
        inter["skip_offsets"] = offs
        return h, inter
    return h


def time_edge_kernel(model, data: Any, flush_buf: Tensor, layer: int = 1, reps: int = 10):
    """CUDA-event time of the fused edge kernel of conv `layer` alone (for bench.py's roofline)."""
    x, n_pulses, ptr, graph, N, B = model._graph_inputs(data, None)
    dev = x.device
    conv = model._conv_layers[layer]
    pk = _conv_pack(conv, _io_dtype(model), first=(layer == 0))
    g = torch.Generator(device=dev).manual_seed(0)
    P = (torch.randn((N, pk.Kp), generator=g, device=dev) * 0.5).to(torch.float16)
    Q = (torch.randn((N, pk.Kp), generator=g, device=dev) * 0.5).to(torch.float16)
    out = torch.empty((N, pk.b), dtype=_io_dtype(model), device=dev)
    coords = torch.empty((N, 3), dtype=torch.float32, device=dev)
    ell = ops.knn_graph_ell(x, conv.nb_neighbors, ptr, [0, 1, 2])
    slots = (_slots_min if model.min_edge_slots else _slots)(conv.nb_neighbors)

    def run():
        ops.tc_edge_mlp_aggregate(P, Q, pk.Kp, ell, slots, pk.W2p, pk.b, out, coords, [0, 1, 2])
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
    E = int(ell.deg.sum().item())
    executed = 2.0 * (N * slots) * pk.Kp * pk.b16
    name = (f"gn::tc::edge_mlp_aggregate_pair_kernel<{slots},false,{'true' if out.dtype == torch.float16 else 'false'},false,false> "
            "(tcgen05.mma.cta_group::2 kind::f16, W2 split over the TMEM of a CTA pair; one launch = "
            "one C-ABI call)")
    return ms, executed, name
