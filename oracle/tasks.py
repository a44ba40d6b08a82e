"""Oracle restatement of the task heads and losses on the path (CPU, torch).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Follows /root/reference/src/graphnet/models/task/task.py:284-295
(`LearnedTask.forward`: affine -> `_forward` -> prediction transform) and
models/task/reconstruction.py:11-112 (the `_forward` of each head), with
`eps_like` from utilities/maths.py:6-8.  Losses follow
training/loss_functions.py:34-60 (`LossFunction.forward`: elements * weights ->
mean), :101-120 (LogCosh) and :354-426,497-520 (von Mises-Fisher 3D); the m=3
normaliser is pinned by the closed form in the reference's own test
tests/training/test_loss_functions.py:71-101.
"""

from __future__ import annotations

import math
from typing import Callable, Dict, Optional

import torch
from torch import Tensor

TASK_NB_INPUTS = {
    "EnergyReconstruction": 1,
    "DirectionReconstructionWithKappa": 3,
    "ZenithReconstruction": 1,
    "ZenithReconstructionWithKappa": 2,
    "AzimuthReconstructionWithKappa": 2,
    "IdentityTask": None,
}


def _eps(x: Tensor) -> float:
    return torch.finfo(x.dtype).eps


def head_transform(kind: str, x: Tensor) -> Tensor:
    """The `_forward` of one head applied to the affine output `x` [B, p]."""
    if kind == "EnergyReconstruction":  # reconstruction.py:109-112
        return torch.nn.functional.softplus(x, beta=0.05) + _eps(x)
    if kind == "DirectionReconstructionWithKappa":  # :63-69
        kappa = torch.linalg.vector_norm(x, dim=1) + _eps(x)
        return torch.stack(
            (x[:, 0] / kappa, x[:, 1] / kappa, x[:, 2] / kappa, kappa), dim=1
        )
    if kind == "ZenithReconstruction":  # :80-82
        return torch.sigmoid(x[:, :1]) * math.pi
    if kind == "ZenithReconstructionWithKappa":  # :94-98
        angle = (torch.sigmoid(x[:, :1]) * math.pi).squeeze(1)
        kappa = torch.abs(x[:, 1]) + _eps(x)
        return torch.stack((angle, kappa), dim=1)
    if kind == "AzimuthReconstructionWithKappa":  # :19-27
        kappa = torch.linalg.vector_norm(x, dim=1) + _eps(x)
        angle = torch.atan2(x[:, 1], x[:, 0])
        angle = torch.where(angle < 0, angle + 2 * math.pi, angle)
        return torch.stack((angle, kappa), dim=1)
    if kind == "IdentityTask":
        return x
    raise ValueError(kind)


def task_forward(
    kind: str,
    weight: Tensor,
    bias: Tensor,
    x: Tensor,
    transform: Optional[Callable[[Tensor], Tensor]] = None,
) -> Tensor:
    """`LearnedTask.forward` (task.py:284-295)."""
    y = torch.nn.functional.linear(x, weight, bias)
    y = head_transform(kind, y)
    return transform(y) if transform is not None else y


# --------------------------------------------------------------------------- #
# losses
# --------------------------------------------------------------------------- #
def _reduce(elements: Tensor, weights: Optional[Tensor]) -> Tensor:
    """LossFunction.forward (loss_functions.py:34-60)."""
    if weights is not None:
        elements = elements * weights
    return torch.mean(elements)


def logcosh_loss(pred: Tensor, target: Tensor, weights: Optional[Tensor] = None) -> Tensor:
    """LogCoshLoss (loss_functions.py:101-120)."""
    diff = pred - target
    el = diff + torch.nn.functional.softplus(-2.0 * diff) - math.log(2.0)
    return _reduce(el, weights)


def log_cmk_m3_exact(kappa: Tensor) -> Tensor:
    """log C_3(kappa) = log k - k - log(2 pi (1 - exp(-2k))).

    Closed form used by the reference's test as ground truth
    (tests/training/test_loss_functions.py:71-101) for the `LogCMK` exact
    branch (loss_functions.py:233-351, m = 3).
    """
    k = kappa.double()
    return (torch.log(k) - k - torch.log(2 * math.pi * (1 - torch.exp(-2 * k)))).to(
        kappa.dtype
    )


def log_cmk_approx(m: int, kappa: Tensor) -> Tensor:
    """loss_functions.py:366-379 (`log_cmk_approx`)."""
    v = m / 2.0 - 0.5
    a = torch.sqrt((v + 1) ** 2 + kappa**2)
    b = v - 1
    return -a + b * torch.log(b + a)


def log_cmk(m: int, kappa: Tensor, kappa_switch: float = 100.0) -> Tensor:
    """loss_functions.py:381-401: exact below `kappa_switch`, approx above."""
    assert m == 3
    ks = torch.tensor([kappa_switch], dtype=kappa.dtype)
    mask_exact = kappa < kappa_switch
    offset = log_cmk_approx(m, ks) - log_cmk_m3_exact(ks)
    ret = log_cmk_approx(m, kappa) - offset
    ret = torch.where(mask_exact, log_cmk_m3_exact(kappa.clamp(max=kappa_switch)), ret)
    return ret


def vmf3d_loss(pred: Tensor, target: Tensor, weights: Optional[Tensor] = None) -> Tensor:
    """VonMisesFisher3DLoss (loss_functions.py:497-520) over `_evaluate` (:403-426)."""
    assert target.dim() == 2 and target.shape[1] == 3
    assert pred.dim() == 2 and pred.shape[1] == 4
    kappa = pred[:, 3]
    p = kappa.unsqueeze(1) * pred[:, [0, 1, 2]]
    k = torch.norm(p, dim=1)
    dotprod = torch.sum(p * target, dim=1)
    el = -log_cmk(3, k) - dotprod
    return _reduce(el, weights)
