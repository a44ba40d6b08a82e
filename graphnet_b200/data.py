"""Minimal `Data` / `Batch` containers with the fields DynEdge consumes.

The reference uses `torch_geometric.data.Data/Batch`.  When
torch_geometric is installed its objects can be passed straight to
`graphnet_b200.DynEdge` - only attribute access (`x`, `edge_index`, `batch`,
`n_pulses`) is used.  These stand-ins exist so that the drop-in also works (and
is testable) where torch_geometric is absent.
"""

from __future__ import annotations

from typing import Any, Dict, Iterable, List

import torch
from torch import Tensor


class Data:
    """Attribute bag: x [n,F], edge_index [2,E], n_pulses, labels ..."""

    def __init__(self, **kwargs: Any) -> None:
        for k, v in kwargs.items():
            setattr(self, k, v)

    def __getitem__(self, key: str) -> Any:
        return getattr(self, key)

    def __setitem__(self, key: str, value: Any) -> None:
        setattr(self, key, value)

    def __contains__(self, key: str) -> bool:
        return hasattr(self, key)

    def keys(self) -> List[str]:
        return [k for k in self.__dict__ if not k.startswith("_")]

    def to(self, device: Any, non_blocking: bool = False) -> "Data":
        out = self.__class__()
        for k, v in self.__dict__.items():
            out.__dict__[k] = v.to(device, non_blocking=non_blocking) if isinstance(v, Tensor) else v
        return out

    @property
    def num_nodes(self) -> int:
        return int(self.x.shape[0])


class Batch(Data):
    """Disjoint-union batch of events (`Batch.from_data_list`)."""

    @classmethod
    def from_data_list(cls, data_list: Iterable[Data]) -> "Batch":
        data_list = list(data_list)
        out = cls()
        sizes = [int(d.x.shape[0]) for d in data_list]
        offsets = [0]
        for s in sizes:
            offsets.append(offsets[-1] + s)
        keys = data_list[0].keys()
        for k in keys:
            vals = [getattr(d, k) for d in data_list]
            v0 = vals[0]
            if k == "edge_index":
                out.edge_index = torch.cat([v + o for v, o in zip(vals, offsets[:-1])], dim=1)
            elif isinstance(v0, Tensor) and v0.dim() >= 1 and v0.shape[0] == sizes[0] and k != "n_pulses":
                out[k] = torch.cat(vals, dim=0)  # node-level
            elif isinstance(v0, Tensor):
                out[k] = torch.stack(vals) if v0.dim() == 0 else torch.cat(vals, dim=0)
            else:
                out[k] = vals
        out.batch = torch.repeat_interleave(
            torch.arange(len(sizes)), torch.tensor(sizes, dtype=torch.long))
        out.ptr = torch.tensor(offsets, dtype=torch.long)
        out.num_graphs = len(sizes)
        return out
