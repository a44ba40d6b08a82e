"""Minimal `Data` / `Batch` containers with the fields DynEdge consumes.

The reference uses `torch_geometric.data.Data/Batch`.  When
torch_geometric is installed its objects can be passed straight to
`graphnet_b200.DynEdge` - only attribute access (`x`, `edge_index`, `batch`,
`n_pulses`) is used.  These stand-ins exist so that the drop-in also works (and
is testable) where torch_geometric is absent.
"""

from __future__ import annotations

from typing import Any, Dict, Iterable, List, Optional

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


class DevicePrefetcher:
    """Iterate over host batches, yielding device batches whose host->device copy overlapped the
    work on the previous batch.

    The reference feeds its models from a `torch.utils.data.DataLoader(pin_memory=...)` whose
    batches Lightning moves to the device synchronously before each step.  Here the copy of batch
    i+1 is issued on a side stream as soon as batch i has been handed out, so it runs under
    batch i's kernels; the consumer's stream waits on the copy's event (no host sync).  Tensors
    should be pinned; the 10 MB of a 1024-event IceCube batch (8 MB of it the int64 `edge_index`
    PyG batches carry) then cost no step time.
    """

    def __init__(self, batches: Iterable[Any], device: Any):
        self._it = iter(batches)
        self._device = torch.device(device)
        self._stream = torch.cuda.Stream(device=self._device)
        self._next = None
        self._preload()

    def _preload(self) -> None:
        try:
            host = next(self._it)
        except StopIteration:
            self._next = None
            return
        with torch.cuda.stream(self._stream):
            dev = host.to(self._device, non_blocking=True)
            event = torch.cuda.Event()
            event.record(self._stream)
        self._next = (dev, event, host)

    def __iter__(self) -> "DevicePrefetcher":
        return self

    def __next__(self) -> Any:
        if self._next is None:
            raise StopIteration
        dev, event, _host = self._next
        cur = torch.cuda.current_stream(self._device)
        cur.wait_event(event)
        for v in dev.__dict__.values():      # the caching allocator must not recycle these
            if isinstance(v, Tensor):        # buffers for the side stream while `cur` uses them
                v.record_stream(cur)
        self._preload()
        return dev


class ResultReader:
    """Device -> host read of per-batch results, awaited ONE batch behind the launches.

    `push(tensors)` copies the tensors of the batch just enqueued into pinned host buffers on the
    current stream (asynchronously) and returns the host copies of the PREVIOUS batch, waiting
    only for that batch's copy event: the host enqueues batch i+1 while the device still works on
    batch i, instead of idling the device for the host's launch time after every blocking read.
    `flush()` returns the last batch.  Two sets of pinned buffers are reused alternately: the
    tensors a call returns stay valid until the next call but one.
    """

    def __init__(self) -> None:
        self._slots: List[Optional[tuple]] = [None, None]   # (host tensors, event)
        self._i = 0
        self._pending: Optional[int] = None

    def _host_like(self, slot: int, tensors: List[Tensor]) -> List[Tensor]:
        cur = self._slots[slot]
        if cur is not None and len(cur[0]) == len(tensors) and all(
                h.shape == t.shape and h.dtype == t.dtype for h, t in zip(cur[0], tensors)):
            return cur[0]
        return [torch.empty(t.shape, dtype=t.dtype, pin_memory=True) for t in tensors]

    def _take(self, slot: int) -> List[Tensor]:
        host, event = self._slots[slot]
        event.synchronize()
        return host

    def push(self, tensors: List[Tensor]) -> Optional[List[Tensor]]:
        slot = self._i & 1
        host = self._host_like(slot, tensors)
        for h, t in zip(host, tensors):
            h.copy_(t, non_blocking=True)
        event = torch.cuda.Event()
        event.record()
        self._slots[slot] = (host, event)
        prev, self._pending = self._pending, slot
        self._i += 1
        return None if prev is None else self._take(prev)

    def flush(self) -> Optional[List[Tensor]]:
        prev, self._pending = self._pending, None
        return None if prev is None else self._take(prev)
