"""CPU: the C-ABI library loads and exports every symbol include/*.h declares."""
import ctypes
import glob
import os
import re

import pytest
import torch

from conftest import ROOT


def _declared_symbols():
    names = set()
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        if h.endswith("_debug.h"):   # exports of the profiling build only
            continue
        text = open(h).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        names |= set(re.findall(r"\b(gn_[a-z0-9_]+)\s*\(", text))
    return sorted(names)


def test_header_declares_symbols():
    syms = _declared_symbols()
    assert "gn_knn_graph" in syms and "gn_edgeconv_f32" in syms and len(syms) >= 12


def test_library_exports_every_declared_symbol():
    from graphnet_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in _declared_symbols():
        assert hasattr(lib, name), f"{name} declared in include/ but not exported"


def test_library_exports_nothing_undeclared():
    """Every gn_* symbol of the product library is declared in include/graphnet_b200.h (no
    diagnostic or leftover exports)."""
    import subprocess
    from graphnet_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True,
                         text=True, check=True).stdout
    exported = {line.split()[-1] for line in out.splitlines()
                if line.split() and line.split()[-1].startswith("gn_")}
    assert exported == set(_declared_symbols()), exported ^ set(_declared_symbols())


def test_binding_covers_every_declared_symbol():
    from graphnet_b200 import _lib
    assert sorted(_lib.SIGNATURES) == _declared_symbols()
    lib = _lib.load()
    assert lib.gn_version() >= 100
    assert lib.gn_scan_scratch_bytes(5000) == 5 * 8


def test_no_cpu_fallback():
    import graphnet_b200 as g
    with pytest.raises(RuntimeError):
        g.knn_graph(torch.randn(10, 3), 4)
    m = g.DynEdge(7).eval()
    d = g.Data(x=torch.randn(6, 7), edge_index=torch.zeros(2, 0, dtype=torch.long),
               batch=torch.zeros(6, dtype=torch.long), n_pulses=torch.tensor([6], dtype=torch.int32))
    with pytest.raises(RuntimeError):
        m(d)


def test_product_never_imports_oracle():
    for path in glob.glob(os.path.join(ROOT, "graphnet_b200", "**", "*.py"), recursive=True):
        src = open(path).read()
        assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), path
