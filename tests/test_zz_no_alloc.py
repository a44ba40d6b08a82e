"""GPU: the allocation-free row-offset entry point against the scratch-taking one.

`gn_ell_row_offsets_noalloc` exists for hosts that cannot call cudaMalloc while running (a
detector-side filter process with a fixed GPU memory budget, or a process whose device memory
is entirely owned by another framework's pool).  It keeps the scan's partial sums in an array
compiled into the library.  This file is collected last on purpose: a run that stops at the
first failure has then already executed every other test.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _offsets_with_scratch(lib, deg):
    N = deg.shape[0]
    out = torch.empty(N + 1, dtype=torch.int64, device=DEV)
    nbytes = lib.gn_scan_scratch_bytes(N)
    scratch = torch.empty(max(nbytes, 8), dtype=torch.uint8, device=DEV)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.gn_ell_row_offsets(C.c_void_p(deg.data_ptr()), N, C.c_void_p(out.data_ptr()),
                                C.c_void_p(scratch.data_ptr()), nbytes, st)
    assert rc == 0, lib.gn_last_error()
    return out


def _offsets_noalloc(lib, deg, out):
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    return lib.gn_ell_row_offsets_noalloc(C.c_void_p(deg.data_ptr()), deg.shape[0],
                                          C.c_void_p(out.data_ptr()), st)


def test_noalloc_row_offsets_match_twice_in_a_row():
    from graphnet_b200 import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(0)
    # two different sizes back to back: the second call reuses the built-in array after the
    # first one left its partial sums there (1 block, then 2443 blocks)
    for N in (700, 2_500_001):
        deg = torch.randint(0, 10, (N,), generator=g, dtype=torch.int32).to(DEV)
        want = _offsets_with_scratch(lib, deg)
        out = torch.full((N + 1,), -1, dtype=torch.int64, device=DEV)
        # the whole vector against an independent cumulative sum (not only the sibling scan)
        ref = torch.zeros(N + 1, dtype=torch.int64, device=DEV)
        ref[1:] = torch.cumsum(deg.to(torch.int64), 0)
        torch.cuda.synchronize()
        # the DEVICE's free memory, as the driver reports it: a cudaMalloc made inside the library
        # would not show in torch's own counters (torch cannot see it), but it shows here
        free_before, _ = torch.cuda.mem_get_info()
        torch_before = torch.cuda.memory_allocated()
        assert _offsets_noalloc(lib, deg, out) == 0, lib.gn_last_error()
        torch.cuda.synchronize()
        free_after, _ = torch.cuda.mem_get_info()
        assert free_after == free_before, "the call changed the device's free memory"
        assert torch.cuda.memory_allocated() == torch_before  # nor anything from torch's pool
        assert torch.equal(out, want)  # integer result: bit-exact
        assert torch.equal(out, ref)
    # and the smaller one again, after the large call
    deg = torch.randint(0, 10, (700,), generator=g, dtype=torch.int32).to(DEV)
    out = torch.empty(701, dtype=torch.int64, device=DEV)
    assert _offsets_noalloc(lib, deg, out) == 0
    assert torch.equal(out, _offsets_with_scratch(lib, deg))


def test_noalloc_rejects_sizes_beyond_the_built_in_array():
    from graphnet_b200 import _lib
    lib = _lib.load()
    deg = torch.zeros(8, dtype=torch.int32, device=DEV)
    out = torch.zeros(9, dtype=torch.int64, device=DEV)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.gn_ell_row_offsets_noalloc(C.c_void_p(deg.data_ptr()), 4194304 + 1,
                                        C.c_void_p(out.data_ptr()), st)
    assert rc == -1
    assert b"exceeds the built-in scratch" in lib.gn_last_error()
    # N == 0 is fine and writes the single terminating offset
    rc = lib.gn_ell_row_offsets_noalloc(None, 0, C.c_void_p(out.data_ptr()), st)
    assert rc == 0 and int(out[0]) == 0
