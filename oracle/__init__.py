"""CPU oracle for the DynEdge hot path.  TEST INFRASTRUCTURE ONLY.

This package restates, in plain CPU PyTorch, the algorithm that the reference
(`graphnet-team/graphnet`, mounted at /root/reference when the oracle was
written) executes for the path

    KNNGraph -> DynEdge (4x DynEdgeConv with dynamic kNN) -> global pooling
             -> post-processing / readout MLP -> task heads.

The arithmetic of that path lives in third-party wheels that are NOT vendored
in the reference and NOT installed in the build container (torch_cluster,
torch_scatter, torch_geometric; all unpinned in the reference's setup.py:56-87).
The restatements in `oracle/ops.py` therefore follow the published semantics of
those libraries and the call sites in the reference.

Parity status
-------------
* In-tree logic (layer construction, forward order, global variables, skip-cat,
  heads) is PINNED: `oracle/refstub/` lets the reference's own
  `graphnet.models.gnn.DynEdge` be imported with the third-party ops replaced by
  `oracle/ops.py`, and `tests/golden/make_golden.py` records its outputs.
* Third-party op semantics (kNN tie order, scatter, EdgeConv) are UNPINNED by
  any reference test or golden vector (the reference has none for this path,
  SURVEY.md section 4): "parity unpinned" at that boundary.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline /
`--impl reference` legs may import this package.  The product
(`graphnet_b200/`) never does and has no CPU fallback.
"""

from .ops import (  # noqa: F401
    knn_graph,
    edge_conv,
    scatter_sum,
    scatter_mean,
    scatter_min,
    scatter_max,
    homophily,
)
from .dynedge import (  # noqa: F401
    DynEdgeSpec,
    init_dynedge_state_dict,
    dynedge_forward,
    global_variables,
)
from .tasks import task_forward, TASK_NB_INPUTS  # noqa: F401
