"""Tensor-level wrappers over the C ABI (device memory and streams via PyTorch).

PyTorch is plumbing here: it owns device memory (outputs and scratch come from
its caching allocator) and the current CUDA stream; all arithmetic happens in
the sm_100a kernels of `csrc/` reached through `include/graphnet_b200.h`.
Every wrapper refuses CPU tensors - there is no CPU path.
"""

from __future__ import annotations

import ctypes as C
import os
import threading
from dataclasses import dataclass
from typing import Any, Callable, List, Optional, Sequence, Tuple, Union

import torch
from torch import Tensor

from . import _lib
from ._lib import (ACT, AGGR, DTYPE_BF16, DTYPE_BF16X3, DTYPE_F16, DTYPE_F32, HEAD, POOL, GnGraph,
                   check)


def _require_cuda(*tensors: Optional[Tensor]) -> torch.device:
    dev = None
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise RuntimeError(
                "graphnet_b200 operates on CUDA tensors only (no CPU fallback); "
                f"got a tensor on {t.device}"
            )
        dev = t.device if dev is None else dev
        if t.device != dev:
            raise RuntimeError("graphnet_b200: tensors on different devices")
    assert dev is not None
    return dev


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)
_capture = threading.local()   # .stream: the library's capture stream while run_captured is active


def _stream(dev: torch.device) -> C.c_void_p:
    # the raw handle of torch's current stream (no Stream object per call where torch offers it)
    cap = getattr(_capture, "stream", None)
    if cap is not None:
        return C.c_void_p(cap)
    if _raw_stream is None:
        return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    return C.c_void_p(_raw_stream(dev.index if dev.index is not None
                                  else torch.cuda.current_device()))


class _NoSwitch:
    def __enter__(self):
        return None

    def __exit__(self, *exc):
        return False


_NO_SWITCH = _NoSwitch()


def _on(dev: torch.device):
    """Context that makes `dev` the current CUDA device - free when it already is."""
    if dev.index is None or dev.index == torch.cuda.current_device():
        return _NO_SWITCH
    return torch.cuda.device(dev)


def run_captured(key: int, dev: torch.device, fn: Callable[[], Any]) -> Any:
    """Run `fn` - a sequence of this module's calls on `dev` and `torch.empty` allocations, NO
    torch kernels - with every launch captured and replayed as one executable CUDA graph on the
    current stream (gn_graph_capture_begin / _launch: the graph cached under `key` is updated
    with this call's parameters, so shapes may change from call to call).  No idle GPU between
    the kernels.  Falls back to running `fn` directly when capturing is not possible (the
    current stream is itself capturing, GN_NO_GRAPH=1, or the driver refuses the capture)."""
    if os.environ.get("GN_NO_GRAPH") or getattr(_capture, "stream", None) is not None or \
            torch.cuda.is_current_stream_capturing():
        return fn()
    lib = _lib.load()
    with _on(dev):
        real = _stream(dev)
        handle = C.c_void_p()
        if lib.gn_graph_capture_begin(C.byref(handle)) != 0:
            return fn()
        _capture.stream = handle.value
        try:
            out = fn()
        except BaseException:
            _capture.stream = None
            lib.gn_graph_capture_abort()
            raise
        _capture.stream = None
        if lib.gn_graph_capture_launch(C.c_void_p(key), real) != 0:
            return fn()      # not capturable here (first use of a kernel): direct launches
    return out


def zeros(shape, dtype: torch.dtype, device: torch.device) -> Tensor:
    """`torch.zeros` whose fill is a launch of this library on the current (or capture) stream -
    inside `run_captured` a torch fill kernel would run before the captured launches do."""
    t = torch.empty(shape, dtype=dtype, device=device)
    with _on(t.device):
        check(_lib.load().gn_zero(C.c_void_p(t.data_ptr()), t.numel() * t.element_size(),
                                  _stream(t.device)), "gn_zero")
    return t


def _ptr(t: Optional[Tensor]) -> C.c_void_p:
    return C.c_void_p(0 if t is None else t.data_ptr())


def _host_i32(vals: Sequence[int]):
    return (C.c_int32 * len(vals))(*[int(v) for v in vals])


@dataclass
class Graph:
    """In-edges grouped by target node (ELL from our kNN, or CSR from user COO)."""

    col: Tensor  # int32
    deg: Optional[Tensor]  # int32 [N]   (ELL)
    rowptr: Optional[Tensor]  # int32 [N+1] (CSR)
    ell_stride: int
    num_nodes: int
    flags: Optional[Tensor] = None  # int32[1], device
    max_degree: Optional[Tensor] = None  # int32[1], device

    def c_struct(self) -> GnGraph:
        return GnGraph(
            col=self.col.data_ptr(),
            deg=0 if self.deg is None else self.deg.data_ptr(),
            rowptr=0 if self.rowptr is None else self.rowptr.data_ptr(),
            ell_stride=self.ell_stride,
            num_edges=0 if self.rowptr is None else int(self.col.shape[0]),
        )


# ------------------------------------------------------------------ graph build
def batch_to_ptr(batch: Tensor, num_events: int) -> Tuple[Tensor, Tensor]:
    """Event offsets `ptr` [B+1] (int32) from the sorted `batch` vector, plus flags."""
    dev = _require_cuda(batch)
    batch = batch.contiguous()
    if batch.dtype != torch.int64:
        batch = batch.long()
    ptr = torch.empty(num_events + 1, dtype=torch.int32, device=dev)
    flags = torch.zeros(1, dtype=torch.int32, device=dev)
    with _on(dev):
        check(_lib.load().gn_batch_to_ptr(_ptr(batch), batch.shape[0], num_events, _ptr(ptr),
                                          _ptr(flags), _stream(dev)), "gn_batch_to_ptr")
    return ptr, flags


def _columns(features_subset: Union[slice, Sequence[int], None], width: int) -> List[int]:
    if features_subset is None:
        return list(range(width))
    if isinstance(features_subset, slice):
        return list(range(width))[features_subset]
    return [int(c) if c >= 0 else width + int(c) for c in features_subset]


def knn_graph_ell(x: Tensor, k: int, ptr: Tensor, columns: Sequence[int], loop: bool = False,
                  stride: Optional[int] = None) -> Graph:
    """kNN graph of `x[:, columns]` per event in ELL layout (device only, no sync)."""
    dev = _require_cuda(x, ptr)
    assert x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1
    N = x.shape[0]
    B = ptr.shape[0] - 1
    kk = k + (0 if loop else 1)
    stride = kk if stride is None else stride
    nbr = torch.empty((N, stride), dtype=torch.int32, device=dev)
    deg = torch.empty(N, dtype=torch.int32, device=dev)
    with _on(dev):
        check(_lib.load().gn_knn_graph(_ptr(x), x.stride(0), _host_i32(columns), len(columns),
                                       _ptr(ptr), N, B, k, int(loop), _ptr(nbr), stride,
                                       _ptr(deg), _stream(dev)), "gn_knn_graph")
    return Graph(col=nbr, deg=deg, rowptr=None, ell_stride=stride, num_nodes=N)


def ell_to_coo(g: Graph) -> Tensor:
    """int64 `edge_index` [2, E] from an ELL graph (one host sync for E)."""
    assert g.ell_stride > 0 and g.deg is not None
    dev = g.col.device
    N = g.num_nodes
    lib = _lib.load()
    rowptr = torch.empty(N + 1, dtype=torch.int64, device=dev)
    nbytes = lib.gn_scan_scratch_bytes(N)
    scratch = torch.empty(max(nbytes, 8), dtype=torch.uint8, device=dev)
    with _on(dev):
        check(lib.gn_ell_row_offsets(_ptr(g.deg), N, _ptr(rowptr), _ptr(scratch), nbytes,
                                     _stream(dev)), "gn_ell_row_offsets")
        E = int(rowptr[N].item())
        ei = torch.empty((2, E), dtype=torch.int64, device=dev)
        check(lib.gn_ell_to_coo(_ptr(g.col), g.ell_stride, _ptr(g.deg), _ptr(rowptr), N, E,
                                _ptr(ei), _stream(dev)), "gn_ell_to_coo")
    return ei


def knn_graph(x: Tensor, k: int, batch: Optional[Tensor] = None, loop: bool = False,
              num_events: Optional[int] = None) -> Tensor:
    """Drop-in for `torch_geometric.nn.knn_graph(x, k, batch, loop)` on CUDA tensors.

    Reference call sites: graphs/edges/edges.py:78-82, components/layers.py:70-74.
    Returns int64 COO [2, E] (row 0 = neighbour j, row 1 = query i), ordered by
    query, then ascending (distance^2, index).
    """
    dev = _require_cuda(x, batch)
    x = x.float().contiguous()
    N = x.shape[0]
    if N == 0:
        return torch.zeros((2, 0), dtype=torch.int64, device=dev)
    if batch is None:
        ptr = torch.tensor([0, N], dtype=torch.int32, device=dev)
    else:
        if num_events is None:
            num_events = int(batch.max().item()) + 1
        ptr, flags = batch_to_ptr(batch, num_events)
        if int(flags.item()) != 0:
            raise ValueError("knn_graph: `batch` must be sorted with ids in [0, num_events)")
    g = knn_graph_ell(x, k, ptr, list(range(x.shape[1])), loop=loop)
    return ell_to_coo(g)


def coo_to_csr(edge_index: Tensor, num_nodes: int) -> Graph:
    """User `edge_index` [2, E] (int64, targets non-decreasing) -> CSR by target."""
    dev = _require_cuda(edge_index)
    assert edge_index.dim() == 2 and edge_index.shape[0] == 2
    ei = edge_index.contiguous()
    if ei.dtype != torch.int64:
        ei = ei.long()
    E = ei.shape[1]
    # zero-filled: for an edge list with ids out of range the kernel flags the error and leaves
    # part of rowptr unwritten - consumers launched before the flags are read must then see empty
    # rows, not garbage offsets
    rowptr = torch.zeros(num_nodes + 1, dtype=torch.int32, device=dev)
    col = torch.empty(max(E, 1), dtype=torch.int32, device=dev)
    flags = torch.zeros(1, dtype=torch.int32, device=dev)
    maxdeg = torch.empty(1, dtype=torch.int32, device=dev)
    with _on(dev):
        check(_lib.load().gn_coo_to_csr(_ptr(ei), E, num_nodes, _ptr(rowptr), _ptr(col),
                                        _ptr(flags), _ptr(maxdeg), _stream(dev)), "gn_coo_to_csr")
    return Graph(col=col, deg=None, rowptr=rowptr, ell_stride=0, num_nodes=num_nodes,
                 flags=flags, max_degree=maxdeg)


# ----------------------------------------------------------- per-event features
_DT = {torch.float32: DTYPE_F32, torch.bfloat16: DTYPE_BF16, torch.float16: DTYPE_F16}


def global_variables(x: Tensor, graph: Graph, ptr: Tensor, n_pulses: Tensor,
                     x0: Optional[Tensor] = None, x0_cols: int = 0,
                     x0b: Optional[Tensor] = None, x0b_cols: int = 0,
                     x0_split: bool = False) -> Tensor:
    """gv [B, F+5]; if `x0` (/`x0b`) is given also writes [:, :cols] = [x, gv[batch], 0...].
    `x0_split`: x0 is a bf16 [N, 3*x0_cols] buffer that receives [hi | lo | hi] (GN_BF16X3)."""
    dev = _require_cuda(x, ptr, n_pulses, x0, x0b)
    assert x.dtype == torch.float32 and x.stride(1) == 1
    N, F = x.shape
    B = ptr.shape[0] - 1
    n_pulses = n_pulses.to(torch.int32).contiguous()
    assert n_pulses.shape[0] == B, "n_pulses must have one entry per event"
    gv = torch.empty((B, F + 5), dtype=torch.float32, device=dev)
    gs = graph.c_struct()
    outs = []
    for t in (x0, x0b):
        if t is not None:
            assert t.stride(1) == 1 and t.shape[0] == N
            outs.append((_DT[t.dtype], t.stride(0)))
        else:
            outs.append((DTYPE_F32, 0))
    if x0_split:
        assert x0 is not None and x0.dtype == torch.bfloat16 and x0.shape[1] >= 3 * x0_cols
        outs[0] = (DTYPE_BF16X3, x0.stride(0))
    with _on(dev):
        check(_lib.load().gn_global_variables(
            _ptr(x), x.stride(0), F, C.byref(gs), _ptr(ptr), _ptr(n_pulses), N, B, _ptr(gv),
            _ptr(x0), outs[0][1], x0_cols, outs[0][0], _ptr(x0b), outs[1][1], x0b_cols,
            outs[1][0], _stream(dev)), "gn_global_variables")
    return gv


def segment_pool(x: Tensor, ptr: Tensor, schemes: Sequence[str],
                 out: Optional[Tensor] = None) -> Tensor:
    """Per-event min/max/sum/mean of `x` [N, C], concatenated per scheme -> [B, C*P]."""
    dev = _require_cuda(x, ptr, out)
    assert x.stride(1) == 1
    C_ = x.shape[1]
    B = ptr.shape[0] - 1
    P = len(schemes)
    if out is None:
        out = torch.empty((B, C_ * P), dtype=torch.float32, device=dev)
    assert out.dtype == torch.float32 and out.stride(1) == 1
    # gn_segment_pool reads float32 or bfloat16 rows (no fp16 loader): anything else would be
    # reinterpreted silently
    assert x.dtype in (torch.float32, torch.bfloat16), f"segment_pool: unsupported dtype {x.dtype}"
    dt = DTYPE_BF16 if x.dtype == torch.bfloat16 else DTYPE_F32
    with _on(dev):
        check(_lib.load().gn_segment_pool(_ptr(x), dt, x.stride(0), C_, _ptr(ptr), B,
                                          _host_i32([POOL[s] for s in schemes]), P, _ptr(out),
                                          out.stride(0), _stream(dev)), "gn_segment_pool")
    return out


# ------------------------------------------------------------------ dense, fp32
def linear_f32(A: Tensor, W: Tensor, bias: Optional[Tensor], act: str = "none",
               out: Optional[Tensor] = None) -> Tensor:
    """act(A W^T + bias) with torch.nn.Linear's weight layout [N, K]."""
    dev = _require_cuda(A, W, bias, out)
    assert A.dtype == torch.float32 and A.stride(1) == 1
    W = W.detach().contiguous()
    M, K = A.shape
    Nout = W.shape[0]
    assert W.shape[1] == K
    if out is None:
        out = torch.empty((M, Nout), dtype=torch.float32, device=dev)
    assert out.stride(1) == 1 and out.shape[0] == M and out.shape[1] >= Nout
    b = None if bias is None else bias.detach().contiguous()
    dt = {torch.bfloat16: DTYPE_BF16, torch.float16: DTYPE_F16, torch.float32: DTYPE_F32}[out.dtype]
    with _on(dev):
        check(_lib.load().gn_linear_f32(_ptr(A), A.stride(0), _ptr(W), _ptr(b), _ptr(out),
                                        out.stride(0), dt, M, Nout, K, ACT[act], _stream(dev)),
              "gn_linear_f32")
    return out


def edgeconv_prepare_f32(W1: Tensor, W2: Optional[Tensor]) -> Tuple[Tensor, Optional[Tensor]]:
    """Repack an EdgeConv MLP's weights: Wpq [2a, d], W2T [a, b]."""
    dev = _require_cuda(W1, W2)
    W1 = W1.detach().float().contiguous()
    a, d2 = W1.shape
    assert d2 % 2 == 0
    d = d2 // 2
    Wpq = torch.empty((2 * a, d), dtype=torch.float32, device=dev)
    if W2 is not None:
        W2 = W2.detach().float().contiguous()
        b = W2.shape[0]
        assert W2.shape[1] == a
        W2T = torch.empty((a, b), dtype=torch.float32, device=dev)
    else:
        b, W2T = a, None
    with _on(dev):
        check(_lib.load().gn_edgeconv_prepare_f32(_ptr(W1), _ptr(W2), d, a, b, _ptr(Wpq),
                                                  _ptr(W2T), _stream(dev)),
              "gn_edgeconv_prepare_f32")
    return Wpq, W2T


def edgeconv_f32(x: Tensor, graph: Graph, Wpq: Tensor, b1: Optional[Tensor],
                 W2T: Optional[Tensor], b2: Optional[Tensor], act: str = "relu",
                 aggr: str = "add", out: Optional[Tensor] = None) -> Tensor:
    """Fused EdgeConv forward (fp32): out[i] = aggr_j act(W2 act(W1 [x_i, x_j-x_i] + b1) + b2)."""
    dev = _require_cuda(x, Wpq, W2T, out)
    assert x.dtype == torch.float32 and x.stride(1) == 1
    N, d = x.shape
    a = Wpq.shape[0] // 2
    assert Wpq.shape[1] == d
    b = a if W2T is None else W2T.shape[1]
    if out is None:
        out = torch.empty((N, b), dtype=torch.float32, device=dev)
    assert out.stride(1) == 1 and out.shape[0] == N and out.shape[1] >= b
    pq = torch.empty((N, 2 * a), dtype=torch.float32, device=dev)
    zeros = None
    if b1 is None:
        b1 = torch.zeros(a, dtype=torch.float32, device=dev)
    if b2 is None and W2T is not None:
        b2 = torch.zeros(b, dtype=torch.float32, device=dev)
    gs = graph.c_struct()
    with _on(dev):
        check(_lib.load().gn_edgeconv_f32(
            _ptr(x), x.stride(0), d, C.byref(gs), N, _ptr(Wpq), _ptr(b1.detach().contiguous()), a,
            _ptr(W2T), _ptr(None if b2 is None else b2.detach().contiguous()), b, ACT[act],
            AGGR[aggr], _ptr(pq), _ptr(out), out.stride(0), _stream(dev)), "gn_edgeconv_f32")
    return out


def task_head(y: Tensor, kind: str) -> Tensor:
    """The `_forward` transform of a reconstruction head on the affine output `y`."""
    dev = _require_cuda(y)
    y = y.float().contiguous()
    M, p = y.shape
    code = HEAD[kind]
    p_out = 4 if code == 2 else (2 if code in (4, 5) else (1 if code == 3 else p))
    out = torch.empty((M, p_out), dtype=torch.float32, device=dev)
    with _on(dev):
        check(_lib.load().gn_task_head(_ptr(y), M, p, code, _ptr(out), _stream(dev)),
              "gn_task_head")
    return out


# ------------------------------------------------- dense, bf16 tensor cores
def tc_linear_bf16(A: Tensor, W: Tensor, bias: Optional[Tensor], out: Tensor, K: int,
                   act: str = "none") -> Tensor:
    """out[:, :N] = act(A[:, :K] W[:, :K]^T + bias); A, W bf16 or fp16; fp32 accumulate (tcgen05)."""
    dev = _require_cuda(A, W, bias, out)
    assert A.dtype == W.dtype and A.dtype in (torch.bfloat16, torch.float16)
    assert A.stride(1) == 1 and W.stride(1) == 1 and out.stride(1) == 1
    M = A.shape[0]
    Nout = W.shape[0]
    assert out.shape[0] == M and out.shape[1] >= Nout and A.shape[1] >= K and W.shape[1] >= K
    dt = {torch.bfloat16: DTYPE_BF16, torch.float16: DTYPE_F16, torch.float32: DTYPE_F32}[out.dtype]
    with _on(dev):
        check(_lib.load().gn_tc_linear(_ptr(A), A.stride(0), _ptr(W), W.stride(0), _DT[A.dtype],
                                       _ptr(bias), _ptr(out), out.stride(0), dt, M, Nout, K,
                                       ACT[act], _stream(dev)), "gn_tc_linear")
    return out


def tc_linear_pool(A: Tensor, W: Tensor, bias: Optional[Tensor], K: int, batch: Tensor,
                   ptr: Tensor, schemes: Sequence[str], out: Tensor,
                   out16: Optional[Tensor] = None, scale16: float = 1.0) -> Tensor:
    """out[e, s*C + c] = pool_s over the rows of event e of relu(A[:, :K] W[:, :K]^T + bias)[:, c]:
    the last post-processing Linear fused with the global pooling (the [N, C] matrix is never
    stored).  `out16` (optional, fp16): the same values times `scale16`."""
    dev = _require_cuda(A, W, bias, batch, ptr, out, out16)
    assert A.dtype == W.dtype and A.dtype in (torch.bfloat16, torch.float16)
    assert A.stride(1) == 1 and W.stride(1) == 1 and out.stride(1) == 1
    assert batch.dtype == torch.int64 and batch.is_contiguous() and ptr.dtype == torch.int32
    M, C_ = A.shape[0], W.shape[0]
    B = ptr.shape[0] - 1
    P = len(schemes)
    assert batch.shape[0] == M and out.dtype == torch.float32 and out.shape[0] == B
    assert out.shape[1] >= C_ * P and A.shape[1] >= K and W.shape[1] >= K
    if out16 is not None:
        assert out16.dtype == torch.float16 and out16.stride(1) == 1 and out16.shape[1] >= C_ * P
    assert (B * C_) % 2 == 0, "tc_linear_pool: B * C must be even (64-bit sum plane)"
    acc = torch.empty((4, B, C_), dtype=torch.float32, device=dev)
    lib = _lib.load()
    with _on(dev):
        check(lib.gn_tc_linear_pool(_ptr(A), A.stride(0), _ptr(W), W.stride(0), _DT[A.dtype],
                                    _ptr(bias), M, C_, K, _ptr(batch), _ptr(ptr), B, _ptr(acc),
                                    _stream(dev)), "gn_tc_linear_pool")
        check(lib.gn_pool_finalize(_ptr(acc), _ptr(ptr), B, C_,
                                   _host_i32([POOL[s] for s in schemes]), P, _ptr(out),
                                   out.stride(0), _ptr(out16),
                                   out16.stride(0) if out16 is not None else 0, float(scale16),
                                   _stream(dev)), "gn_pool_finalize")
    return out


def tc_edge_mlp_aggregate(P: Tensor, Q: Tensor, Kp: int, graph: Graph, slots: int, W2p: Tensor,
                          b: int, out: Tensor, coords: Optional[Tensor],
                          coord_cols: Sequence[int], relu_mask: Optional[Tensor] = None,
                          act: str = "relu", aggr: str = "add") -> Tensor:
    """out[i] = aggr_j act(W2 act(P_i + Q_j) + b2) on tcgen05 (see include/graphnet_b200.h);
    act in relu / gelu / leaky_relu, aggr in add / mean / max / min.

    `relu_mask` ([8, ld] int32 from `edge_relu_mask`, training): receives the ReLU bits of the
    second Linear for the backward pass (gn_tc_edge_mlp_aggregate_mask)."""
    dev = _require_cuda(P, Q, W2p, out, coords, relu_mask)
    assert P.dtype == torch.float16 and Q.dtype == torch.float16 and W2p.dtype == torch.float16
    assert out.dtype in (torch.bfloat16, torch.float16)
    assert P.stride(1) == 1 and Q.stride(1) == 1 and P.shape[0] == Q.shape[0]
    N = P.shape[0]
    lib = _lib.load()
    gs = graph.c_struct()
    nbytes = lib.gn_tc_edge_scratch_bytes(C.byref(gs), N, slots)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    ncoord = 0 if coords is None else len(coord_cols)
    with _on(dev):
        if relu_mask is not None:
            assert relu_mask.dtype == torch.int32 and relu_mask.shape[0] == 8 \
                and relu_mask.stride(1) == 1
            check(lib.gn_tc_edge_mlp_aggregate_mask(
                _ptr(P), P.stride(0), _ptr(Q), Q.stride(0), Kp, C.byref(gs), N, slots, _ptr(W2p),
                b, _ptr(out), out.stride(0), _DT[out.dtype], _ptr(coords),
                _host_i32(coord_cols) if ncoord else None, ncoord, _ptr(scratch), nbytes,
                _ptr(relu_mask), relu_mask.stride(0), _stream(dev)),
                "gn_tc_edge_mlp_aggregate_mask")
            return out
        check(lib.gn_tc_edge_mlp_aggregate(
            _ptr(P), P.stride(0), _ptr(Q), Q.stride(0), Kp, C.byref(gs), N, slots, _ptr(W2p), b,
            _ptr(out), out.stride(0), _DT[out.dtype], _ptr(coords),
            _host_i32(coord_cols) if ncoord else None,
            ncoord, _ptr(scratch), nbytes, ACT[act], AGGR[aggr], _stream(dev)),
            "gn_tc_edge_mlp_aggregate")
    return out


def tc_edge_mlp_aggregate_f32(P: Tensor, Q: Tensor, Kp: int, graph: Graph, slots: int, W2hi: Tensor,
                              W2lo: Tensor, b: int, out: Tensor, coords: Optional[Tensor],
                              coord_cols: Sequence[int]) -> Tensor:
    """out[i] = sum_j relu(W2 relu(P_i + Q_j) + b2) with fp32 P, Q, out and (hi, lo) fp16 operand
    pairs on the tensor cores (gn_tc_edge_mlp_aggregate_f32): the fp32 path's EdgeConv."""
    dev = _require_cuda(P, Q, W2hi, W2lo, out, coords)
    assert P.dtype == torch.float32 and Q.dtype == torch.float32 and out.dtype == torch.float32
    assert W2hi.dtype == torch.float16 and W2lo.dtype == torch.float16
    assert W2hi.is_contiguous() and W2lo.is_contiguous() and W2hi.shape == W2lo.shape
    assert P.stride(1) == 1 and Q.stride(1) == 1 and out.stride(1) == 1 and P.shape[0] == Q.shape[0]
    N = P.shape[0]
    lib = _lib.load()
    gs = graph.c_struct()
    nbytes = lib.gn_tc_edge_scratch_bytes(C.byref(gs), N, slots)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    ncoord = 0 if coords is None else len(coord_cols)
    with _on(dev):
        check(lib.gn_tc_edge_mlp_aggregate_f32(
            _ptr(P), P.stride(0), _ptr(Q), Q.stride(0), Kp, C.byref(gs), N, slots, _ptr(W2hi),
            _ptr(W2lo), b, _ptr(out), out.stride(0), _ptr(coords),
            _host_i32(coord_cols) if ncoord else None, ncoord, _ptr(scratch), nbytes,
            _stream(dev)), "gn_tc_edge_mlp_aggregate_f32")
    return out


def split_bf16(X: Tensor, kc: int, terms: int = 3, out: Optional[Tensor] = None) -> Tensor:
    """bf16 pieces of the fp32 matrix X [M, cols <= kc] for a split-operand GEMM (gn_split_bf16):
    terms=3 -> [x1 | x2 | x1], terms=6 -> [x1 | x1 | x1 | x2 | x2 | x3], blocks kc wide."""
    dev = _require_cuda(X, out)
    assert X.dtype == torch.float32 and X.stride(1) == 1 and kc % 4 == 0 and kc >= X.shape[1]
    M, cols = X.shape
    if out is None:
        out = torch.empty((M, terms * kc), dtype=torch.bfloat16, device=dev)
    assert out.dtype == torch.bfloat16 and out.stride(1) == 1 and out.shape[1] >= terms * kc
    with _on(dev):
        check(_lib.load().gn_split_bf16(_ptr(X), X.stride(0), M, cols, kc, terms, _ptr(out),
                                        out.stride(0), _stream(dev)), "gn_split_bf16")
    return out


def split_bf16_both(X: Tensor, kc: int, out3: Tensor, out6: Optional[Tensor] = None) -> Tensor:
    """Three-term pieces of X into `out3` and six-term pieces into `out6` (returned) from one read."""
    dev = _require_cuda(X, out3, out6)
    assert X.dtype == torch.float32 and X.stride(1) == 1 and kc % 4 == 0 and kc >= X.shape[1]
    M, cols = X.shape
    if out6 is None:
        out6 = torch.empty((M, 6 * kc), dtype=torch.bfloat16, device=dev)
    for o, t in ((out3, 3), (out6, 6)):
        assert o.dtype == torch.bfloat16 and o.stride(1) == 1 and o.shape[1] >= t * kc
    with _on(dev):
        check(_lib.load().gn_split_bf16_both(_ptr(X), X.stride(0), M, cols, kc, _ptr(out3),
                                             out3.stride(0), _ptr(out6), out6.stride(0),
                                             _stream(dev)), "gn_split_bf16_both")
    return out6


def edgeconv_pack_tc(W1: Tensor, b1: Optional[Tensor], W2: Tensor, b2: Optional[Tensor], Kp: int,
                     Kd: int, wpq_dtype: torch.dtype, train: bool, act: str = "relu"):
    """(Wpq, bias_pq, W2p, W2b, W2bT, WpqT) for the tensor-core EdgeConv kernels from the
    nn.Linear parameters, in one launch; the last three only when `train` (else None)."""
    dev = _require_cuda(W1, b1, W2, b2)
    W1, W2 = W1.detach().contiguous(), W2.detach().contiguous()
    b1 = None if b1 is None else b1.detach().contiguous()
    b2 = None if b2 is None else b2.detach().contiguous()
    assert W1.dtype == torch.float32 and W2.dtype == torch.float32
    a, d2 = W1.shape
    d, b = d2 // 2, W2.shape[0]
    b16 = (b + 15) // 16 * 16
    Wpq = torch.empty((2 * Kp, Kd), dtype=wpq_dtype, device=dev)
    bias = torch.empty(2 * Kp, dtype=torch.float32, device=dev)
    W2p = torch.empty((b16, Kp), dtype=torch.float16, device=dev)
    W2b = W2bT = WpqT = None
    if train:
        W2b = torch.empty((b16, Kp), dtype=torch.bfloat16, device=dev)
        W2bT = torch.empty((Kp, b16), dtype=torch.bfloat16, device=dev)
        if wpq_dtype != torch.float32:
            WpqT = torch.empty((Kd, 2 * Kp), dtype=wpq_dtype, device=dev)
    with _on(dev):
        check(_lib.load().gn_edgeconv_pack_tc(_ptr(W1), _ptr(b1), _ptr(W2), _ptr(b2), d, a, b, Kp,
                                              Kd, _ptr(Wpq), _DT[wpq_dtype], _ptr(bias), _ptr(W2p),
                                              _ptr(W2b), _ptr(W2bT), _ptr(WpqT), ACT[act],
                                              _stream(dev)),
              "gn_edgeconv_pack_tc")
    return Wpq, bias, W2p, W2b, W2bT, WpqT


def edge_relu_mask(N: int, slots: int, device) -> Tensor:
    """Buffer for the ReLU bits of one EdgeConv layer: [8, ld] int32 (8 groups of 32 channels)."""
    ld = int(_lib.load().gn_tc_edge_relu_mask_ld(N, slots))
    if ld <= 0:
        raise ValueError(f"no ReLU-bit layout for {slots} slots per node")
    return torch.empty((8, ld), dtype=torch.int32, device=device)


# ------------------------------------------------- backward pass (16-bit training path)
def _dt_of(t: Tensor) -> int:
    return {torch.bfloat16: DTYPE_BF16, torch.float16: DTYPE_F16, torch.float32: DTYPE_F32}[t.dtype]


def tc_linear_accumulate(A: Tensor, W: Tensor, out: Tensor, K: int, splits: int = 0) -> Tensor:
    """out[:, :N] += A[:, :K] W[:, :K]^T (fp32 red.add; split K when few output tiles)."""
    dev = _require_cuda(A, W, out)
    assert A.dtype == W.dtype and A.dtype in (torch.bfloat16, torch.float16)
    assert out.dtype == torch.float32
    assert A.stride(1) == 1 and W.stride(1) == 1 and out.stride(1) == 1
    M, Nout = A.shape[0], W.shape[0]
    assert out.shape[0] == M and out.shape[1] >= Nout
    with _on(dev):
        check(_lib.load().gn_tc_linear_accumulate(_ptr(A), A.stride(0), _ptr(W), W.stride(0),
                                                  _DT[A.dtype], _ptr(out), out.stride(0), M, Nout,
                                                  K, splits, _stream(dev)),
              "gn_tc_linear_accumulate")
    return out


def transpose_16(src: Tensor, out: Optional[Tensor] = None) -> Tensor:
    """[rows, cols] 16-bit -> [cols, rows padded to a multiple of 8] (row stride of the result)."""
    dev = _require_cuda(src, out)
    assert src.dtype in (torch.bfloat16, torch.float16) and src.stride(1) == 1
    rows, cols = src.shape
    if out is None:
        ld = (rows + 7) // 8 * 8
        out = torch.empty((cols, ld), dtype=src.dtype, device=dev)
    assert out.stride(1) == 1 and out.shape[0] >= cols and out.shape[1] >= rows
    with _on(dev):
        check(_lib.load().gn_transpose_16(_ptr(src), src.stride(0), rows, cols, _ptr(out),
                                          out.stride(0), _stream(dev)), "gn_transpose_16")
    return out


def colsum_accumulate(X: Tensor, out: Tensor) -> Tensor:
    """out[n] += sum_m X[m, n] (fp32)."""
    dev = _require_cuda(X, out)
    assert X.stride(1) == 1 and out.dtype == torch.float32 and out.is_contiguous()
    M, N = X.shape
    assert out.numel() >= N
    with _on(dev):
        check(_lib.load().gn_colsum_accumulate(_ptr(X), _dt_of(X), X.stride(0), M, N, _ptr(out),
                                               _stream(dev)), "gn_colsum_accumulate")
    return out


def relu_backward(dY: Tensor, Y: Optional[Tensor], out: Optional[Tensor] = None) -> Tensor:
    """bf16(dY * [Y > 0]) (Y None: plain conversion)."""
    dev = _require_cuda(dY, Y, out)
    M, N = dY.shape
    if out is None:
        out = torch.empty((M, (N + 7) // 8 * 8), dtype=torch.bfloat16, device=dev)[:, :N]
    assert out.dtype == torch.bfloat16 and dY.stride(1) == 1 and out.stride(1) == 1
    assert Y is None or (Y.stride(1) == 1 and Y.shape[0] == M and Y.shape[1] >= N)
    with _on(dev):
        check(_lib.load().gn_relu_backward(_ptr(dY), _dt_of(dY), dY.stride(0), _ptr(Y),
                                           0 if Y is None else _dt_of(Y),
                                           0 if Y is None else Y.stride(0), _ptr(out),
                                           out.stride(0), M, N, _stream(dev)), "gn_relu_backward")
    return out


def edge_build_h(P: Tensor, Q: Tensor, Kp: int, graph: Graph, H: Optional[Tensor],
                 HT: Optional[Tensor]) -> None:
    dev = _require_cuda(P, Q, H, HT)
    assert P.dtype == torch.float16 and Q.dtype == torch.float16
    assert (H is None or H.dtype == torch.bfloat16) and (H is not None or HT is not None)
    N = P.shape[0]
    gs = graph.c_struct()
    with _on(dev):
        check(_lib.load().gn_edge_build_h(_ptr(P), P.stride(0), _ptr(Q), Q.stride(0), Kp,
                                          C.byref(gs), N, _ptr(H),
                                          0 if H is None else H.stride(0), _ptr(HT),
                                          0 if HT is None else HT.stride(0), _stream(dev)),
              "gn_edge_build_h")


def edge_grad_z(Z: Tensor, dOut: Tensor, b: int, graph: Graph, G2: Tensor, G2T: Tensor) -> None:
    dev = _require_cuda(Z, dOut, G2, G2T)
    assert Z.dtype == torch.bfloat16 and dOut.dtype == torch.float32
    assert G2.dtype == torch.bfloat16 and G2T.dtype == torch.bfloat16
    N = dOut.shape[0]
    gs = graph.c_struct()
    with _on(dev):
        check(_lib.load().gn_edge_grad_z(_ptr(Z), Z.stride(0), _ptr(dOut), dOut.stride(0), b,
                                         C.byref(gs), N, _ptr(G2), G2.stride(0), _ptr(G2T),
                                         G2T.stride(0), _stream(dev)), "gn_edge_grad_z")


def edge_grad_z_mask(relu_mask: Tensor, dOut: Tensor, b: int, graph: Graph, G2: Tensor,
                     G2T: Tensor) -> None:
    """G2 / G2T from the ReLU bits the training forward recorded (no Z)."""
    dev = _require_cuda(relu_mask, dOut, G2, G2T)
    assert relu_mask.dtype == torch.int32 and dOut.dtype == torch.float32
    assert G2.dtype == torch.bfloat16 and G2T.dtype == torch.bfloat16
    N = dOut.shape[0]
    gs = graph.c_struct()
    with _on(dev):
        check(_lib.load().gn_edge_grad_z_mask(_ptr(relu_mask), relu_mask.stride(0), _ptr(dOut),
                                              dOut.stride(0), b, C.byref(gs), N, _ptr(G2),
                                              G2.stride(0), _ptr(G2T), G2T.stride(0),
                                              _stream(dev)), "gn_edge_grad_z_mask")


def ell_reverse(graph: Graph, ptr: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
    """(rstart [N], rcount [N], rslot [N*stride]): the slots of an ELL graph grouped by source."""
    dev = _require_cuda(graph.col, ptr)
    N, S = graph.num_nodes, graph.ell_stride
    assert S > 0 and ptr.dtype == torch.int32
    rstart = torch.empty(N, dtype=torch.int32, device=dev)
    rcount = torch.empty(N, dtype=torch.int32, device=dev)
    rslot = torch.empty(max(N * S, 1), dtype=torch.int32, device=dev)
    gs = graph.c_struct()
    with _on(dev):
        check(_lib.load().gn_ell_reverse(C.byref(gs), _ptr(ptr), N, ptr.shape[0] - 1, _ptr(rstart),
                                         _ptr(rcount), _ptr(rslot), _stream(dev)), "gn_ell_reverse")
    return rstart, rcount, rslot


def edge_grad_pq_gather(dH: Tensor, P: Tensor, Q: Tensor, Kp: int, graph: Graph, ptr: Tensor,
                        reverse: Optional[Tuple[Tensor, Tensor, Tensor]], dP: Tensor, dQ: Tensor,
                        dP32: Optional[Tensor] = None, dQ32: Optional[Tensor] = None) -> None:
    """dP / dQ (bf16, fp32 copies) from dH.  `reverse` None: one pass, dQ accumulated in `dQ32`
    (required) by vector reductions; `ell_reverse(...)`: list form, fixed summation order."""
    dev = _require_cuda(dH, P, Q, dP, dQ, dP32, dQ32)
    assert dH.dtype == torch.bfloat16 and P.dtype == torch.float16 and Q.dtype == torch.float16
    assert dP.dtype == torch.bfloat16 and dQ.dtype == torch.bfloat16
    N = dP.shape[0]
    rstart, rcount, rslot = reverse if reverse is not None else (None, None, None)
    gs = graph.c_struct()
    with _on(dev):
        check(_lib.load().gn_edge_grad_pq_gather(
            _ptr(dH), dH.stride(0), _ptr(P), P.stride(0), _ptr(Q), Q.stride(0), Kp, C.byref(gs),
            _ptr(ptr), N, ptr.shape[0] - 1, _ptr(rstart), _ptr(rcount), _ptr(rslot),
            _ptr(dP), dP.stride(0), _ptr(dQ), dQ.stride(0),
            _ptr(dP32), 0 if dP32 is None else dP32.stride(0),
            _ptr(dQ32), 0 if dQ32 is None else dQ32.stride(0), _stream(dev)),
            "gn_edge_grad_pq_gather")


def edge_grad_pq(dH: Tensor, H: Tensor, Kp: int, graph: Graph, dP: Tensor, dQ: Tensor,
                 dP32: Optional[Tensor] = None) -> None:
    dev = _require_cuda(dH, H, dP, dQ, dP32)
    assert dH.dtype == torch.bfloat16 and H.dtype == torch.bfloat16
    assert dP.dtype == torch.bfloat16 and dQ.dtype == torch.float32
    N = dP.shape[0]
    gs = graph.c_struct()
    with _on(dev):
        check(_lib.load().gn_edge_grad_pq(_ptr(dH), dH.stride(0), _ptr(H), H.stride(0), Kp,
                                          C.byref(gs), N, _ptr(dP), dP.stride(0), _ptr(dQ),
                                          dQ.stride(0), _ptr(dP32),
                                          0 if dP32 is None else dP32.stride(0), _stream(dev)),
              "gn_edge_grad_pq")


def outer_accumulate_f32(G: Tensor, X: Tensor, out: Tensor) -> Tensor:
    """out[c, k] += sum_n G[n, c] X[n, k] (fp32, X at most 64 columns wide)."""
    dev = _require_cuda(G, X, out)
    assert G.dtype == X.dtype == out.dtype == torch.float32
    assert G.stride(1) == 1 and X.stride(1) == 1 and out.stride(1) == 1
    N, cols = G.shape
    K = X.shape[1]
    assert X.shape[0] == N and out.shape[0] >= cols and out.shape[1] >= K
    with _on(dev):
        check(_lib.load().gn_outer_accumulate_f32(_ptr(G), G.stride(0), cols, _ptr(X), X.stride(0),
                                                  K, N, _ptr(out), out.stride(0), _stream(dev)),
              "gn_outer_accumulate_f32")
    return out


def segment_pool_backward(x: Tensor, ptr: Tensor, schemes: Sequence[str], pooled: Tensor,
                          d_pooled: Tensor) -> Tensor:
    """Gradient of `segment_pool` with respect to x -> [N, C] fp32."""
    dev = _require_cuda(x, ptr, pooled, d_pooled)
    assert x.stride(1) == 1 and pooled.dtype == torch.float32 and d_pooled.dtype == torch.float32
    assert pooled.stride(1) == 1 and d_pooled.stride(1) == 1
    N, C_ = x.shape
    B = ptr.shape[0] - 1
    dX = zeros((N, C_), torch.float32, dev)
    # gn_segment_pool reads float32 or bfloat16 rows (no fp16 loader): anything else would be
    # reinterpreted silently
    assert x.dtype in (torch.float32, torch.bfloat16), f"segment_pool: unsupported dtype {x.dtype}"
    dt = DTYPE_BF16 if x.dtype == torch.bfloat16 else DTYPE_F32
    with _on(dev):
        check(_lib.load().gn_segment_pool_backward(
            _ptr(x), dt, x.stride(0), C_, _ptr(ptr), B, _host_i32([POOL[s] for s in schemes]),
            len(schemes), _ptr(pooled), pooled.stride(0), _ptr(d_pooled), d_pooled.stride(0),
            _ptr(dX), dX.stride(0), _stream(dev)), "gn_segment_pool_backward")
    return dX


def csr_to_ell(graph: Graph, stride: int) -> Graph:
    """Slot-table (ELL) form of a CSR graph; `stride` must be >= its maximum in-degree."""
    if graph.ell_stride > 0:
        return graph
    dev = _require_cuda(graph.col, graph.rowptr)
    N = graph.num_nodes
    nbr = torch.empty((N, stride), dtype=torch.int32, device=dev)
    deg = torch.empty(N, dtype=torch.int32, device=dev)
    flags = graph.flags if graph.flags is not None else torch.zeros(1, dtype=torch.int32, device=dev)
    with _on(dev):
        check(_lib.load().gn_csr_to_ell(_ptr(graph.rowptr), _ptr(graph.col), N, stride, _ptr(nbr),
                                        _ptr(deg), _ptr(flags), _stream(dev)), "gn_csr_to_ell")
    return Graph(col=nbr, deg=deg, rowptr=None, ell_stride=stride, num_nodes=N, flags=flags,
                 max_degree=graph.max_degree)


# ------------------------------------------------- add_norm_layer=True (fp32, unfused)
def layernorm_act_(X: Tensor, gamma: Optional[Tensor], beta: Optional[Tensor], eps: float,
                   act: str = "none") -> Tensor:
    """In place: X <- act(LayerNorm(X) * gamma + beta), row-wise (torch.nn.LayerNorm semantics)."""
    dev = _require_cuda(X, gamma, beta)
    assert X.dtype == torch.float32 and X.stride(1) == 1
    M, N = X.shape
    g = None if gamma is None else gamma.detach().float().contiguous()
    b = None if beta is None else beta.detach().float().contiguous()
    with _on(dev):
        check(_lib.load().gn_layernorm_act_f32(_ptr(X), X.stride(0), M, N, _ptr(g), _ptr(b),
                                               float(eps), ACT[act], _stream(dev)),
              "gn_layernorm_act_f32")
    return X


def edge_preact_f32(pq: Tensor, a: int, graph: Graph) -> Tensor:
    """[N*stride, a] rows P_i + Q_j of every slot of an ELL graph (zero rows for empty slots)."""
    dev = _require_cuda(pq)
    assert pq.dtype == torch.float32 and pq.is_contiguous() and pq.shape[1] == 2 * a
    N = pq.shape[0]
    out = torch.empty((N * graph.ell_stride, a), dtype=torch.float32, device=dev)
    gs = graph.c_struct()
    with _on(dev):
        check(_lib.load().gn_edge_preact_f32(_ptr(pq), a, C.byref(gs), N, _ptr(out), out.stride(0),
                                             _stream(dev)), "gn_edge_preact_f32")
    return out


def slot_sum_f32(Z: Tensor, graph: Graph, out: Tensor) -> Tensor:
    """out[i] = sum of Z[i*stride + t] over the filled slots of node i."""
    dev = _require_cuda(Z, out)
    assert Z.dtype == torch.float32 and out.dtype == torch.float32
    assert Z.stride(1) == 1 and out.stride(1) == 1
    N, b = out.shape[0], Z.shape[1]
    gs = graph.c_struct()
    with _on(dev):
        check(_lib.load().gn_slot_sum_f32(_ptr(Z), Z.stride(0), b, C.byref(gs), N, _ptr(out),
                                          out.stride(0), _stream(dev)), "gn_slot_sum_f32")
    return out


def slot_aggregate_f32(Z: Tensor, graph: Graph, out: Tensor, aggr: str) -> Tensor:
    """out[i] = add / mean / max / min of Z[i*stride + t] over the filled slots of node i."""
    dev = _require_cuda(Z, out)
    assert Z.dtype == torch.float32 and out.dtype == torch.float32
    assert Z.stride(1) == 1 and out.stride(1) == 1
    N, b = out.shape[0], Z.shape[1]
    gs = graph.c_struct()
    with _on(dev):
        check(_lib.load().gn_slot_aggregate_f32(_ptr(Z), Z.stride(0), b, C.byref(gs), N, AGGR[aggr],
                                                _ptr(out), out.stride(0), _stream(dev)),
              "gn_slot_aggregate_f32")
    return out


def convert_scale_16(X: Tensor, dtype: torch.dtype, scale: float = 1.0) -> Tensor:
    """16-bit copy of an fp32 matrix, multiplied by `scale` (a power of two keeps it exact)."""
    dev = _require_cuda(X)
    assert X.dtype == torch.float32 and X.stride(1) == 1
    M, N = X.shape
    out = torch.empty((M, (N + 7) // 8 * 8), dtype=dtype, device=dev)[:, :N]
    with _on(dev):
        check(_lib.load().gn_convert_scale_16(_ptr(X), X.stride(0), _ptr(out), _DT[dtype],
                                              out.stride(0), M, N, float(scale), _stream(dev)),
              "gn_convert_scale_16")
    return out
