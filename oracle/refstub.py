"""Import the reference's OWN hot-path modules with third-party ops stubbed.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Works only where
/root/reference exists (the build container); nothing in the `-m gpu` tests,
`smoke()` or `bench.py` uses it.  It is how `tests/golden/make_golden.py`
produces the committed golden vectors.

What it does
------------
The reference cannot be imported as a package here: `graphnet.models.__init__`
pulls in pytorch_lightning, torch_geometric, torch_cluster, torch_scatter,
torch_sparse, colorlog, ruamel.yaml ... none of which is installed.  But the
files that hold the hot-path logic are plain Python:

    src/graphnet/models/gnn/dynedge.py          (DynEdge)
    src/graphnet/models/components/layers.py    (DynEdgeConv)
    src/graphnet/models/utils.py                (calculate_xyzt_homophily)
    src/graphnet/models/gnn/gnn.py              (GNN base)
    src/graphnet/models/task/{task,reconstruction}.py
    src/graphnet/training/loss_functions.py
    src/graphnet/utilities/maths.py

`load_reference()` registers empty namespace packages for `graphnet`,
`graphnet.models`, ... (so no `__init__.py` of the reference runs), registers
stand-in modules for the absent third-party packages whose few used entry
points forward to `oracle/ops.py`, and then executes those reference files
unmodified from where they lie.  The result: the reference's own forward code
runs, with only the library kernels replaced by the oracle's restatement.
"""

from __future__ import annotations

import importlib
import os
import sys
import types
from typing import Any, Optional

import torch

from . import ops

REF_SRC = "/root/reference/src"


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REF_SRC, "graphnet"))


class _Placeholder:
    """Stands in for any third-party name the hot path never touches."""

    def __init__(self, *a: Any, **k: Any) -> None:
        pass

    def __class_getitem__(cls, item: Any) -> Any:
        return cls


class _StubModule(types.ModuleType):
    def __getattr__(self, name: str) -> Any:
        if name.startswith("__"):
            raise AttributeError(name)
        return _Placeholder


def _stub(name: str, **attrs: Any) -> types.ModuleType:
    m = _StubModule(name)
    m.__path__ = []  # behave like a package so submodule imports resolve
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules[name] = m
    parent, _, child = name.rpartition(".")
    if parent and parent in sys.modules:
        setattr(sys.modules[parent], child, m)
    return m


class _Data:
    """Attribute bag standing in for torch_geometric.data.Data."""

    def __init__(self, **kw: Any) -> None:
        for k, v in kw.items():
            setattr(self, k, v)

    def __getitem__(self, k: str) -> Any:
        return getattr(self, k)

    def __setitem__(self, k: str, v: Any) -> None:
        setattr(self, k, v)


class _LightningModule(torch.nn.Module):
    @property
    def device(self) -> torch.device:
        for p in self.parameters():
            return p.device
        return torch.device("cpu")


class _MessagePassing(torch.nn.Module):
    def __init__(self, aggr: str = "add", **kw: Any) -> None:
        super().__init__()
        self.aggr = aggr


class _EdgeConv(_MessagePassing):
    """torch_geometric.nn.EdgeConv -> oracle.ops.edge_conv."""

    def __init__(self, nn: Any, aggr: str = "max", **kw: Any) -> None:
        super().__init__(aggr=aggr)
        self.nn = nn

    def forward(self, x: torch.Tensor, edge_index: torch.Tensor) -> torch.Tensor:
        return ops.edge_conv(x, edge_index, self.nn, aggr=self.aggr)


def _scatter_fn(fn: Any) -> Any:
    def wrapped(src, index, dim=0, out=None, dim_size=None):  # noqa: ANN001
        assert dim == 0 and out is None
        return fn(src, index, dim_size)

    return wrapped


def _knn_graph(x, k, batch=None, loop=False, **kw):  # noqa: ANN001
    return ops.knn_graph(x, k, batch, loop=loop)


class _Model(_LightningModule):
    """Minimal stand-in for graphnet.models.Model (logging/config stripped)."""

    def warning(self, *a: Any, **k: Any) -> None:
        pass

    info = debug = error = warning_once = warning


def load_reference() -> types.SimpleNamespace:
    """Return a namespace with the reference's own DynEdge, tasks and losses."""
    if not reference_available():
        raise RuntimeError("/root/reference is not present on this machine")

    # third-party stand-ins -------------------------------------------------
    _stub("pytorch_lightning", LightningModule=_LightningModule)
    _stub("torch_scatter",
          scatter_sum=_scatter_fn(ops.scatter_sum),
          scatter_add=_scatter_fn(ops.scatter_sum),
          scatter_mean=_scatter_fn(ops.scatter_mean),
          scatter_min=_scatter_fn(ops.scatter_min),
          scatter_max=_scatter_fn(ops.scatter_max))
    _stub("torch_sparse")
    _stub("torch_cluster", knn_graph=_knn_graph)
    _stub("torch_geometric")
    _stub("torch_geometric.nn", EdgeConv=_EdgeConv, knn_graph=_knn_graph)
    _stub("torch_geometric.nn.pool", knn_graph=_knn_graph)
    _stub("torch_geometric.nn.conv", MessagePassing=_MessagePassing)
    _stub("torch_geometric.nn.inits")
    _stub("torch_geometric.typing")
    _stub("torch_geometric.data", Data=_Data, Batch=_Data)
    _stub("torch_geometric.utils", homophily=ops.homophily)
    _stub("torch_geometric.utils.num_nodes")

    # reference packages as bare namespaces (no __init__.py executed) --------
    def ns(name: str) -> types.ModuleType:
        m = types.ModuleType(name)
        m.__path__ = [os.path.join(REF_SRC, *name.split("."))]
        sys.modules[name] = m
        parent, _, child = name.rpartition(".")
        if parent:
            setattr(sys.modules[parent], child, m)
        return m

    ns("graphnet")
    models = ns("graphnet.models")
    models.Model = _Model
    ns("graphnet.models.gnn")
    ns("graphnet.models.components")
    ns("graphnet.models.task")
    ns("graphnet.training")
    ns("graphnet.utilities")
    _stub("graphnet.models.model", Model=_Model)
    _stub("graphnet.utilities.imports", has_jammy_flows_package=lambda: False)

    dynedge = importlib.import_module("graphnet.models.gnn.dynedge")
    jinst = importlib.import_module("graphnet.models.gnn.dynedge_jinst")
    particlenet = importlib.import_module("graphnet.models.gnn.particlenet")
    layers = importlib.import_module("graphnet.models.components.layers")
    task = importlib.import_module("graphnet.models.task.task")
    # `from graphnet.models.task import StandardLearnedTask` in reconstruction.py
    sys.modules["graphnet.models.task"].StandardLearnedTask = task.StandardLearnedTask
    recon = importlib.import_module("graphnet.models.task.reconstruction")
    loss = importlib.import_module("graphnet.training.loss_functions")
    return types.SimpleNamespace(
        DynEdge=dynedge.DynEdge,
        DynEdgeJINST=jinst.DynEdgeJINST,
        ParticleNeT=particlenet.ParticleNeT,
        DynEdgeConv=layers.DynEdgeConv,
        Data=_Data,
        task=task,
        reconstruction=recon,
        loss_functions=loss,
    )
