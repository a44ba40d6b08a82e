"""graphnet_b200: the DynEdge hot path of GraphNeT on B200 (sm_100a) kernels.

Host side mirrors the reference's interface for this path (`DynEdge`,
`DynEdgeConv`, `KNNGraph`, `KNNEdges`, `knn_graph`, `StandardModel`, task
heads); compute goes through the C ABI in include/graphnet_b200.h.  Importing
the package does not need the CUDA library; calling any compute entry point
does, and fails loudly when it is missing (no CPU fallback).
"""

from .data import Batch, Data  # noqa: F401
from .detector import IceCube86, IceCubeUpgrade, Prometheus, ORCA150SuperDense  # noqa: F401
from .dynedge import DynEdge, DynEdgeConv  # noqa: F401
from .dynedge_jinst import DynEdgeJINST  # noqa: F401
from .particlenet import ParticleNeT  # noqa: F401
from .graphs import KNNEdges, KNNGraph  # noqa: F401
from .losses import LogCoshLoss, VonMisesFisher3DLoss  # noqa: F401
from .ops import knn_graph  # noqa: F401
from .standard_model import StandardModel  # noqa: F401
from . import training  # noqa: F401
from .graphed import GraphedInference  # noqa: F401
from .tasks import (  # noqa: F401
    AzimuthReconstructionWithKappa,
    DirectionReconstructionWithKappa,
    EnergyReconstruction,
    IdentityTask,
    StandardLearnedTask,
    ZenithReconstruction,
    ZenithReconstructionWithKappa,
)

__version__ = "0.1.0"
