"""Loss functions used by the DynEdge training configs, device-native.

Mirrors /root/reference/src/graphnet/training/loss_functions.py:
`LossFunction.forward` :34-60 (elements * weights -> mean), `LogCoshLoss`
:101-120 and `VonMisesFisher3DLoss` :497-520.  The reference evaluates the vMF
normaliser through `scipy.special.iv` on the HOST every step
(:233-351); for m = 3 the normaliser has the closed form
log C_3(k) = log k - k - log(2 pi (1 - exp(-2k))) - the expression the
reference's own test uses as ground truth
(tests/training/test_loss_functions.py:71-101) - so it is evaluated on the
device here, with no host round trip.
"""

from __future__ import annotations

import math
from typing import Optional

import torch
from torch import Tensor


class LossFunction(torch.nn.Module):
    def forward(self, prediction: Tensor, target: Tensor, weights: Optional[Tensor] = None,
                return_elements: bool = False) -> Tensor:
        elements = self._forward(prediction, target)
        if weights is not None:
            elements = elements * weights
        assert elements.size(dim=0) == target.size(dim=0), "`_forward` should return elementwise loss terms."
        return elements if return_elements else torch.mean(elements)

    def _forward(self, prediction: Tensor, target: Tensor) -> Tensor:
        raise NotImplementedError


class LogCoshLoss(LossFunction):
    def _forward(self, prediction: Tensor, target: Tensor) -> Tensor:
        diff = prediction - target
        return diff + torch.nn.functional.softplus(-2.0 * diff) - math.log(2.0)


class VonMisesFisher3DLoss(LossFunction):
    @staticmethod
    def log_cmk_exact(kappa: Tensor) -> Tensor:
        k = kappa.double()
        return (torch.log(k) - k - torch.log(2 * math.pi * (-torch.expm1(-2 * k)))).to(kappa.dtype)

    @staticmethod
    def log_cmk_approx(kappa: Tensor) -> Tensor:
        v = 3 / 2.0 - 0.5
        a = torch.sqrt((v + 1) ** 2 + kappa**2)
        b = v - 1
        return -a + b * torch.log(b + a)

    @classmethod
    def log_cmk(cls, kappa: Tensor, kappa_switch: float = 100.0) -> Tensor:
        ks = torch.tensor([kappa_switch], dtype=kappa.dtype, device=kappa.device)
        offset = cls.log_cmk_approx(ks) - cls.log_cmk_exact(ks)
        approx = cls.log_cmk_approx(kappa) - offset
        exact = cls.log_cmk_exact(kappa.clamp(max=kappa_switch))
        return torch.where(kappa < kappa_switch, exact, approx)

    def _forward(self, prediction: Tensor, target: Tensor) -> Tensor:
        target = target.reshape(-1, 3)
        assert prediction.dim() == 2 and prediction.size()[1] == 4
        assert prediction.size()[0] == target.size()[0]
        kappa = prediction[:, 3]
        p = kappa.unsqueeze(1) * prediction[:, [0, 1, 2]]
        k = torch.norm(p, dim=1)
        dotprod = torch.sum(p * target, dim=1)
        return -self.log_cmk(k) - dotprod
