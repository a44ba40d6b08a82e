"""Task heads with the reference's interface, running on the CUDA kernels.

Mirrors /root/reference/src/graphnet/models/task/task.py:28-348 (`Task`,
`LearnedTask`, `StandardLearnedTask`) and task/reconstruction.py:11-112 (the
heads).  Parameter name `_affine` is kept so checkpoints load
(`_tasks.{i}._affine.*`).
"""

from __future__ import annotations

from typing import Any, Callable, Dict, List, Optional, Tuple, Union

import numpy as np
import torch
from torch import Tensor

from . import ops


def _head_autograd(x: Tensor, kind: str) -> Tensor:
    """The `_forward` transforms of task/reconstruction.py:11-112 as differentiable torch ops
    (same arithmetic as csrc/heads.cu; used only while autograd is recording)."""
    eps = torch.finfo(x.dtype).eps
    if kind == "IdentityTask":
        return x
    if kind == "EnergyReconstruction":
        return torch.nn.functional.softplus(x, beta=0.05) + eps
    if kind == "DirectionReconstructionWithKappa":
        kappa = torch.linalg.vector_norm(x, dim=1) + eps
        return torch.cat([x[:, :3] / kappa[:, None], kappa[:, None]], dim=1)
    if kind == "ZenithReconstruction":
        return torch.sigmoid(x[:, :1]) * np.pi
    if kind == "ZenithReconstructionWithKappa":
        return torch.stack([torch.sigmoid(x[:, 0]) * np.pi, x[:, 1].abs() + eps], dim=1)
    if kind == "AzimuthReconstructionWithKappa":
        kappa = torch.linalg.vector_norm(x, dim=1) + eps
        angle = torch.atan2(x[:, 1], x[:, 0])
        angle = torch.where(angle < 0, angle + 2 * np.pi, angle)
        return torch.stack([angle, kappa], dim=1)
    raise NotImplementedError(f"training through the {kind} head")


class StandardLearnedTask(torch.nn.Module):
    """affine -> `_forward` head transform -> prediction transform (task.py:284-295)."""

    default_target_labels: List[str] = []
    default_prediction_labels: List[str] = []
    nb_inputs: int = 1
    _head_kind: str = "IdentityTask"

    def __init__(
        self,
        hidden_size: int,
        loss_function: Optional[Callable] = None,
        disable_affine: bool = False,
        *,
        target_labels: Optional[Union[str, List[str]]] = None,
        prediction_labels: Optional[Union[str, List[str]]] = None,
        transform_prediction_and_target: Optional[Callable] = None,
        transform_target: Optional[Callable] = None,
        transform_inference: Optional[Callable] = None,
        transform_support: Optional[Tuple] = None,
        loss_weight: Optional[str] = None,
    ):
        super().__init__()
        if target_labels is None:
            target_labels = self.default_target_labels
        if isinstance(target_labels, str):
            target_labels = [target_labels]
        if prediction_labels is None:
            prediction_labels = self.default_prediction_labels
        if isinstance(prediction_labels, str):
            prediction_labels = [prediction_labels]
        self._target_labels = target_labels
        self._prediction_labels = prediction_labels
        self._inference = False
        self._loss_weight = loss_weight
        self._loss_function = loss_function
        self._regularisation_loss: float = 0.0
        self._disable_affine = disable_affine
        self._affine: torch.nn.Module = (
            torch.nn.Identity() if disable_affine else torch.nn.Linear(hidden_size, self.nb_inputs)
        )
        self._transform_prediction_training: Callable[[Tensor], Tensor] = lambda x: x
        self._transform_prediction_inference: Callable[[Tensor], Tensor] = lambda x: x
        self._transform_target: Callable[[Tensor], Tensor] = lambda x: x
        self._validate_and_set_transforms(
            transform_prediction_and_target, transform_target, transform_inference,
            transform_support)

    # task.py:141-223
    def _validate_and_set_transforms(self, both, target, inference, support) -> None:
        assert not (both is not None and target is not None), (
            "Please specify at most one of `transform_prediction_and_target` and "
            "`transform_target`")
        if target is not None:
            assert inference is not None
            if support is not None:
                assert len(support) == 2, "Please specify min and max for transformation support."
                x_test = torch.from_numpy(np.linspace(support[0], support[1], 10))
            else:
                x_test = np.logspace(-6, 6, 12 + 1)
                x_test = torch.from_numpy(np.concatenate([-x_test[::-1], [0], x_test]))
            try:
                t_test = torch.unsqueeze(target(x_test), -1)
                t_test = torch.squeeze(inference(t_test), -1)
                valid = torch.isfinite(t_test)
                assert torch.allclose(t_test[valid], x_test[valid]), (
                    "The provided transforms for targets during training and predictions "
                    "during inference are not inverse. Please adjust transformation functions "
                    "or support.")
            except IndexError:
                pass
        if both is not None:
            self._transform_prediction_training = both
            self._transform_target = both
        else:
            if target is not None:
                self._transform_target = target
            if inference is not None:
                self._transform_prediction_inference = inference

    @property
    def default_target_labels_(self) -> List[str]:
        return self.default_target_labels

    def inference(self) -> None:
        """Activate inference mode (task.py:131-134)."""
        self._inference = True

    def train_eval(self) -> None:
        """Deactivate inference mode (task.py:136-139)."""
        self._inference = False

    def _transform_prediction(self, prediction: Tensor) -> Tensor:
        if self._inference:
            return self._transform_prediction_inference(prediction)
        return self._transform_prediction_training(prediction)

    def forward(self, x: Tensor) -> Tensor:
        self._regularisation_loss = 0
        # eval mode (model.train(False) / model.eval()) always takes the kernel path below, whether
        # or not the caller wrapped the call in torch.no_grad(); the autograd branch is for training
        if self.training and torch.is_grad_enabled() and (x.requires_grad or any(
                p.requires_grad for p in self._affine.parameters())):
            # training: [B, hidden] -> [B, <=4]; autograd records these few element-wise torch
            # ops (the backbone's backward is the hand-written one, training.py)
            x = self._affine(x.float())
            return self._transform_prediction(_head_autograd(x, self._head_kind))
        with torch.no_grad():
            if isinstance(self._affine, torch.nn.Linear):
                x = ops.linear_f32(x.float().contiguous(), self._affine.weight, self._affine.bias)
            x = ops.task_head(x, self._head_kind)
            return self._transform_prediction(x)

    def compute_loss(self, pred: Tensor, data: Any) -> Tensor:
        """task.py:330-348."""
        target = torch.stack([data[label] for label in self._target_labels], dim=1)
        target = self._transform_target(target)
        weights = data[self._loss_weight] if self._loss_weight is not None else None
        return self._loss_function(pred, target, weights=weights) + self._regularisation_loss


class IdentityTask(StandardLearnedTask):
    """task.py `IdentityTask`: affine only."""

    _head_kind = "IdentityTask"

    def __init__(self, nb_outputs: int, target_labels: Union[List[str], Any], *args: Any,
                 **kwargs: Any):
        self.nb_inputs = nb_outputs  # read by the base ctor when it builds the affine
        self.default_target_labels = (
            target_labels if isinstance(target_labels, list) else [target_labels])
        self.default_prediction_labels = [f"target_{i}_pred" for i in range(nb_outputs)]
        super().__init__(*args, **kwargs)


class EnergyReconstruction(StandardLearnedTask):
    """reconstruction.py:101-112: softplus(x, beta=0.05) + eps."""

    default_target_labels = ["energy"]
    default_prediction_labels = ["energy_pred"]
    nb_inputs = 1
    _head_kind = "EnergyReconstruction"


class DirectionReconstructionWithKappa(StandardLearnedTask):
    """reconstruction.py:49-69: unit vector and kappa = |x| + eps."""

    default_target_labels = ["direction"]
    default_prediction_labels = ["dir_x_pred", "dir_y_pred", "dir_z_pred", "direction_kappa"]
    nb_inputs = 3
    _head_kind = "DirectionReconstructionWithKappa"


class ZenithReconstruction(StandardLearnedTask):
    """reconstruction.py:72-82: sigmoid(x) * pi."""

    default_target_labels = ["zenith"]
    default_prediction_labels = ["zenith_pred"]
    nb_inputs = 1
    _head_kind = "ZenithReconstruction"


class ZenithReconstructionWithKappa(StandardLearnedTask):
    """reconstruction.py:85-98."""

    default_target_labels = ["zenith"]
    default_prediction_labels = ["zenith_pred", "zenith_kappa"]
    nb_inputs = 2
    _head_kind = "ZenithReconstructionWithKappa"


class AzimuthReconstructionWithKappa(StandardLearnedTask):
    """reconstruction.py:11-27."""

    default_target_labels = ["azimuth"]
    default_prediction_labels = ["azimuth_pred", "azimuth_kappa"]
    nb_inputs = 2
    _head_kind = "AzimuthReconstructionWithKappa"
