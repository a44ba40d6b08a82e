"""ctypes binding of the C ABI declared in include/graphnet_b200.h.

The shared library is built in-tree by `__graft_entry__.build()` (plain nvcc,
sm_100a).  There is NO fallback: if the library is missing, importing any
compute entry point raises, and every wrapper refuses non-CUDA tensors.
"""

from __future__ import annotations

import ctypes as C
import os
from typing import Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libgraphnet_b200.so")
if os.environ.get("GN_B200_DEBUG_LIB"):   # profiling build (__graft_entry__.build_debug)
    LIB_PATH = os.path.join(_HERE, "csrc", "libgraphnet_b200_dbg.so")
if os.environ.get("GN_B200_LIB"):         # an explicitly named build (A/B measurements)
    LIB_PATH = os.environ["GN_B200_LIB"]

GN_OK = 0
ACT = {"none": 0, "relu": 1, "gelu": 2, "leaky_relu": 3}
AGGR = {"add": 0, "sum": 0, "mean": 1, "max": 2, "min": 3}
POOL = {"min": 0, "max": 1, "sum": 2, "mean": 3}
DTYPE_F32, DTYPE_BF16, DTYPE_F16, DTYPE_BF16X3 = 0, 1, 2, 3
HEAD = {
    "IdentityTask": 0,
    "EnergyReconstruction": 1,
    "DirectionReconstructionWithKappa": 2,
    "ZenithReconstruction": 3,
    "ZenithReconstructionWithKappa": 4,
    "AzimuthReconstructionWithKappa": 5,
}
FLAG_UNSORTED_BATCH, FLAG_UNSORTED_TARGETS, FLAG_INDEX_OUT_OF_RANGE = 1, 2, 4


class GnGraph(C.Structure):
    _fields_ = [
        ("col", C.c_void_p),
        ("deg", C.c_void_p),
        ("rowptr", C.c_void_p),
        ("ell_stride", C.c_int32),
        ("num_edges", C.c_int64),
    ]


GN_MAX_LAYERS, GN_MAX_DENSE = 8, 4


class GnDynedgeConv(C.Structure):
    _fields_ = [
        ("Wpq", C.c_void_p), ("bias_pq", C.c_void_p), ("W2p", C.c_void_p),
        ("d_in", C.c_int32), ("Kp", C.c_int32), ("b", C.c_int32), ("k", C.c_int32),
        ("slots", C.c_int32), ("act", C.c_int32), ("off_in", C.c_int32), ("off_out", C.c_int32),
        ("ncoord", C.c_int32), ("ldw", C.c_int32), ("coord_cols", C.c_int32 * 16),
    ]


class GnDynedgeDense(C.Structure):
    _fields_ = [
        ("W", C.c_void_p), ("bias", C.c_void_p),
        ("n", C.c_int32), ("K", C.c_int32), ("act", C.c_int32), ("ldw", C.c_int32),
        ("tc", C.c_int32), ("pad_", C.c_int32),
    ]


class GnDynedgePlan(C.Structure):
    _fields_ = [
        ("F", C.c_int32), ("x0w", C.c_int32), ("skip_cols", C.c_int32), ("io_dtype", C.c_int32),
        ("n_conv", C.c_int32), ("n_post", C.c_int32), ("n_readout", C.c_int32),
        ("n_pools", C.c_int32), ("input_k", C.c_int32), ("pad_", C.c_int32),
        ("conv", GnDynedgeConv * GN_MAX_LAYERS),
        ("post", GnDynedgeDense * GN_MAX_DENSE),
        ("readout", GnDynedgeDense * GN_MAX_DENSE),
        ("pools", C.c_int32 * 8),
    ]


_p, _i64, _i32, _int, _sz = C.c_void_p, C.c_int64, C.c_int32, C.c_int, C.c_size_t
_pp = C.POINTER(GnDynedgePlan)
_gp = C.POINTER(GnGraph)
_hp = C.POINTER(C.c_int32)  # host int32 array

# name -> (restype, argtypes); must list every symbol of include/graphnet_b200.h
SIGNATURES = {
    "gn_last_error": (C.c_char_p, []),
    "gn_version": (_int, []),
    "gn_launch_count": (C.c_uint64, []),
    "gn_batch_to_ptr": (_int, [_p, _i64, _i64, _p, _p, _p]),
    "gn_knn_graph": (_int, [_p, _i64, _hp, _int, _p, _i64, _i64, _int, _int, _p, _i32, _p, _p]),
    "gn_scan_scratch_bytes": (_sz, [_i64]),
    "gn_ell_row_offsets": (_int, [_p, _i64, _p, _p, _sz, _p]),
    "gn_ell_row_offsets_noalloc": (_int, [_p, _i64, _p, _p]),
    "gn_ell_to_coo": (_int, [_p, _i32, _p, _p, _i64, _i64, _p, _p]),
    "gn_coo_to_csr": (_int, [_p, _i64, _i64, _p, _p, _p, _p, _p]),
    "gn_global_variables": (_int, [_p, _i64, _int, _gp, _p, _p, _i64, _i64, _p, _p, _i64, _int, _int,
                                   _p, _i64, _int, _int, _p]),
    "gn_segment_pool": (_int, [_p, _int, _i64, _int, _p, _i64, _hp, _int, _p, _i64, _p]),
    "gn_dynedge_infer16_workspace": (_sz, [_pp, _i64, _i64, _i64, _int]),
    "gn_dynedge_infer16": (_int, [_pp, _p, _i64, _p, _p, _i64, _i64, _p, _i64, _p, _sz, _p, _i64, _p, _p]),
    "gn_tc_edge_mlp_aggregate_f32": (_int, [_p, _i64, _p, _i64, _int, _gp, _i64, _int, _p, _p, _int, _p,
                                            _i64, _p, _hp, _int, _p, _sz, _p]),
    "gn_split_bf16": (_int, [_p, _i64, _i64, _int, _int, _int, _p, _i64, _p]),
    "gn_split_bf16_both": (_int, [_p, _i64, _i64, _int, _int, _p, _i64, _p, _i64, _p]),
    "gn_zero": (_int, [_p, _sz, _p]),
    "gn_graph_capture_begin": (_int, [C.POINTER(C.c_void_p)]),
    "gn_graph_capture_launch": (_int, [_p, _p]),
    "gn_graph_capture_abort": (_int, []),
    "gn_tc_linear_pool": (_int, [_p, _i64, _p, _i64, _int, _p, _i64, _int, _int, _p, _p, _i64, _p, _p]),
    "gn_pool_finalize": (_int, [_p, _p, _i64, _int, _hp, _int, _p, _i64, _p, _i64, C.c_float, _p]),
    "gn_linear_f32": (_int, [_p, _i64, _p, _p, _p, _i64, _int, _i64, _int, _int, _int, _p]),
    "gn_edgeconv_prepare_f32": (_int, [_p, _p, _int, _int, _int, _p, _p, _p]),
    "gn_edgeconv_f32": (_int, [_p, _i64, _int, _gp, _i64, _p, _p, _int, _p, _p, _int, _int, _int, _p, _p, _i64, _p]),
    "gn_tc_linear": (_int, [_p, _i64, _p, _i64, _int, _p, _p, _i64, _int, _i64, _int, _int, _int, _p]),
    "gn_convert_scale_16": (_int, [_p, _i64, _p, _int, _i64, _i64, _int, C.c_float, _p]),
    "gn_layernorm_act_f32": (_int, [_p, _i64, _i64, _int, _p, _p, C.c_float, _int, _p]),
    "gn_edge_preact_f32": (_int, [_p, _int, _gp, _i64, _p, _i64, _p]),
    "gn_slot_sum_f32": (_int, [_p, _i64, _int, _gp, _i64, _p, _i64, _p]),
    "gn_slot_aggregate_f32": (_int, [_p, _i64, _int, _gp, _i64, _int, _p, _i64, _p]),
    "gn_tc_linear_accumulate": (_int, [_p, _i64, _p, _i64, _int, _p, _i64, _i64, _int, _int, _int, _p]),
    "gn_transpose_16": (_int, [_p, _i64, _i64, _int, _p, _i64, _p]),
    "gn_csr_to_ell": (_int, [_p, _p, _i64, _int, _p, _p, _p, _p]),
    "gn_colsum_accumulate": (_int, [_p, _int, _i64, _i64, _int, _p, _p]),
    "gn_relu_backward": (_int, [_p, _int, _i64, _p, _int, _i64, _p, _i64, _i64, _int, _p]),
    "gn_edge_build_h": (_int, [_p, _i64, _p, _i64, _int, _gp, _i64, _p, _i64, _p, _i64, _p]),
    "gn_edge_grad_z": (_int, [_p, _i64, _p, _i64, _int, _gp, _i64, _p, _i64, _p, _i64, _p]),
    "gn_edge_grad_pq": (_int, [_p, _i64, _p, _i64, _int, _gp, _i64, _p, _i64, _p, _i64, _p, _i64, _p]),
    "gn_outer_accumulate_f32": (_int, [_p, _i64, _int, _p, _i64, _int, _i64, _p, _i64, _p]),
    "gn_segment_pool_backward": (_int, [_p, _int, _i64, _int, _p, _i64, _p, _int, _p, _i64, _p, _i64, _p, _i64, _p]),
    "gn_tc_edge_scratch_bytes": (_sz, [_gp, _i64, _int]),
    "gn_tc_edge_mlp_aggregate": (_int, [_p, _i64, _p, _i64, _int, _gp, _i64, _int, _p, _int, _p, _i64, _int,
                                        _p, _hp, _int, _p, _sz, _int, _int, _p]),
    "gn_tc_edge_relu_mask_ld": (_i64, [_i64, _int]),
    "gn_tc_edge_mlp_aggregate_mask": (_int, [_p, _i64, _p, _i64, _int, _gp, _i64, _int, _p, _int, _p, _i64,
                                             _int, _p, _hp, _int, _p, _sz, _p, _i64, _p]),
    "gn_edge_grad_z_mask": (_int, [_p, _i64, _p, _i64, _int, _gp, _i64, _p, _i64, _p, _i64, _p]),
    "gn_ell_reverse": (_int, [_gp, _p, _i64, _i64, _p, _p, _p, _p]),
    "gn_edge_grad_pq_gather": (_int, [_p, _i64, _p, _i64, _p, _i64, _int, _gp, _p, _i64, _i64, _p, _p, _p,
                                      _p, _i64, _p, _i64, _p, _i64, _p, _i64, _p]),
    "gn_edgeconv_pack_tc": (_int, [_p, _p, _p, _p, _int, _int, _int, _int, _int, _p, _int, _p, _p, _p, _p, _p,
                                   _int, _p]),
    "gn_task_head": (_int, [_p, _i64, _int, _int, _p, _p]),
}

_lib: Optional[C.CDLL] = None


class GraphnetB200Error(RuntimeError):
    """Raised when a C-ABI call returns a non-zero status."""


def load() -> C.CDLL:
    """Load (once) and return the shared library; raise loudly if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"graphnet_b200: CUDA library not built ({LIB_PATH} missing). "
            "Run `python -c 'import __graft_entry__ as g; g.build()'` in the repo root. "
            "There is no CPU fallback."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(status: int, what: str) -> None:
    if status != GN_OK:
        msg = load().gn_last_error().decode("utf-8", "replace")
        raise GraphnetB200Error(f"{what} failed with status {status}: {msg}")
