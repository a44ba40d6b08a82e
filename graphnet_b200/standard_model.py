"""`StandardModel`: data representation + backbone + task heads.

Mirrors /root/reference/src/graphnet/models/standard_model.py:18-159
(constructor argument names incl. the deprecated `graph_definition` / `gnn`
aliases, `forward(Data | List[Data]) -> List[Tensor]`, `compute_loss`,
`shared_step`) and the `inference()` / `train(mode)` toggling of
easy_model.py:276-289.  The Lightning trainer glue (`fit`, `predict`) is out of
scope; `predict` here is a plain loop over an iterable of batches.
"""

from __future__ import annotations

from typing import Any, Dict, Iterable, List, Optional, Type, Union

import torch
from torch import Tensor

from .tasks import StandardLearnedTask


class StandardModel(torch.nn.Module):
    def __init__(
        self,
        tasks: Union[StandardLearnedTask, List[StandardLearnedTask]],
        data_representation: Any = None,
        graph_definition: Any = None,
        backbone: Optional[torch.nn.Module] = None,
        gnn: Optional[torch.nn.Module] = None,
        optimizer_class: Type[torch.optim.Optimizer] = torch.optim.Adam,
        optimizer_kwargs: Optional[Dict] = None,
        scheduler_class: Optional[type] = None,
        scheduler_kwargs: Optional[Dict] = None,
        scheduler_config: Optional[Dict] = None,
    ) -> None:
        super().__init__()
        if not isinstance(tasks, (list, tuple)):
            tasks = [tasks]
        if data_representation is None and graph_definition is not None:
            data_representation = graph_definition
        elif data_representation is None and graph_definition is None:
            raise TypeError("__init__() missing 1 required keyword argument:'data_representation'")
        if backbone is None and gnn is not None:
            backbone = gnn
        elif backbone is None and gnn is None:
            raise TypeError("__init__() missing 1 required keyword argument:'backbone'")
        assert isinstance(backbone, torch.nn.Module)
        self._tasks = torch.nn.ModuleList(tasks)
        self._data_representation = data_representation
        self.backbone = backbone
        self._optimizer_class = optimizer_class
        self._optimizer_kwargs = optimizer_kwargs or dict()
        self._scheduler_class = scheduler_class
        self._scheduler_kwargs = scheduler_kwargs or dict()
        self._scheduler_config = scheduler_config or dict()
        for task in self._tasks:
            assert isinstance(task, StandardLearnedTask)

    @property
    def target_labels(self) -> List[str]:
        return [label for task in self._tasks for label in task._target_labels]

    @property
    def prediction_labels(self) -> List[str]:
        return [label for task in self._tasks for label in task._prediction_labels]

    def forward(self, data: Any) -> List[Tensor]:
        """standard_model.py:119-132."""
        if not isinstance(data, (list, tuple)):
            data = [data]
        x_list = [self.backbone(d) for d in data]
        x = x_list[0] if len(x_list) == 1 else torch.cat(x_list, dim=0)
        return [task(x) for task in self._tasks]

    def compute_loss(self, preds: List[Tensor], data: Any, verbose: bool = False) -> Tensor:
        """standard_model.py:94-117."""
        if not isinstance(data, (list, tuple)):
            data = [data]
        data_merged: Dict[str, Tensor] = {}
        for label in list(set(self.target_labels)):
            data_merged[label] = torch.cat([d[label] for d in data], dim=0)
        for task in self._tasks:
            if task._loss_weight is not None:
                data_merged[task._loss_weight] = torch.cat(
                    [d[task._loss_weight] for d in data], dim=0)
        losses = [task.compute_loss(pred, data_merged) for task, pred in zip(self._tasks, preds)]
        assert all(loss.dim() == 0 for loss in losses), "Please reduce loss for each task separately"
        return torch.sum(torch.stack(losses))

    def shared_step(self, batch: Any, batch_idx: int = 0) -> Tensor:
        preds = self(batch)
        return self.compute_loss(preds, batch)

    def training_step(self, train_batch: Any, batch_idx: int = 0) -> Tensor:
        """easy_model.py:223-239 without the Lightning logging: the loss of one batch, with
        autograd recording (call `.backward()` on it; the backbone's backward is hand-written,
        training.py)."""
        if isinstance(train_batch, (list, tuple)) and len(train_batch) == 1:
            train_batch = train_batch[0]
        return self.shared_step(train_batch, batch_idx)

    def configure_optimizers(self) -> Any:
        """easy_model.py:77-96: optimizer (+ optional scheduler dict) from the ctor arguments."""
        optimizer = self._optimizer_class(self.parameters(), **self._optimizer_kwargs)
        if self._scheduler_class is None:
            return {"optimizer": optimizer}
        scheduler = self._scheduler_class(optimizer, **self._scheduler_kwargs)
        config = {"optimizer": optimizer, "lr_scheduler": {"scheduler": scheduler}}
        config["lr_scheduler"].update(self._scheduler_config)
        return config

    def inference(self) -> None:
        """easy_model.py:276-279."""
        for task in self._tasks:
            task.inference()

    def train(self, mode: bool = True) -> "StandardModel":
        """easy_model.py:281-289."""
        super().train(mode)
        if mode:
            for task in self._tasks:
                task.train_eval()
        return self

    @torch.no_grad()
    def predict(self, batches: Iterable[Any]) -> List[Tensor]:
        """easy_model.py:291-321 without the Lightning trainer: cat per task."""
        self.inference()
        self.train(False)
        per_batch = [self(b) for b in batches]
        return [torch.cat([pb[i] for pb in per_batch], dim=0) for i in range(len(self._tasks))]

    def predict_to_host(self, host_batches: Iterable[Any], device: Any) -> List[Tensor]:
        """`predict` for a stream of HOST batches (pinned memory), results on the host: the copy
        of batch i+1 runs under batch i's kernels (`DevicePrefetcher`) and the read of batch i's
        predictions is awaited after batch i+1 has been enqueued (`ResultReader`) - the device
        never waits for the host between batches.  Same values as `predict`."""
        from .data import DevicePrefetcher, ResultReader
        self.inference()
        self.train(False)
        reader = ResultReader()
        outs: List[List[Tensor]] = []
        with torch.no_grad():
            for b in DevicePrefetcher(host_batches, device):
                done = reader.push(self(b))
                if done is not None:
                    outs.append([t.clone() for t in done])
            done = reader.flush()
            if done is not None:
                outs.append([t.clone() for t in done])
        return [torch.cat([o[i] for o in outs], dim=0) for i in range(len(self._tasks))]
