"""Event sharding for multi-GPU inference (one process per GPU, `torch.distributed`).

Events are independent (PyG disjoint-union batching; kNN, EdgeConv and pooling never cross
`batch` ids - SURVEY.md 8e), so inference needs no data-path collective: every rank runs the
model on a contiguous range of events and the per-event outputs are gathered at the end.
The reference leaves multi-GPU prediction to Lightning's DistributedSampler
(models/easy_model.py:291-321); this module is the explicit equivalent.
"""

from __future__ import annotations

from typing import Any, Callable, List, Sequence, Tuple

import torch
import torch.distributed as dist
from torch import Tensor

from .data import Data


def shard_ranges(n_pulses: Tensor, world_size: int, balance: str = "pulses") -> List[Tuple[int, int]]:
    """Contiguous event ranges [lo, hi) per rank with balanced cost.

    balance="pulses": cost = n (EdgeConv / GEMM work is linear in nodes);
    balance="pairs":  cost = n^2 (kNN distance work, matters for >1000 pulses/event);
    balance="events": equal counts.
    """
    B = int(n_pulses.shape[0])
    n = n_pulses.to(torch.float64).cpu()
    cost = {"pulses": n, "pairs": n * n, "events": torch.ones_like(n)}[balance]
    csum = torch.cat([torch.zeros(1, dtype=torch.float64), cost.cumsum(0)])
    total = float(csum[-1])
    bounds = [0]
    for r in range(1, world_size):
        target = total * r / world_size
        idx = int(torch.searchsorted(csum, torch.tensor(target, dtype=torch.float64)))
        idx = min(idx, B)
        if idx > 0 and abs(float(csum[idx - 1]) - target) < abs(float(csum[idx]) - target):
            idx -= 1  # the closer of the two neighbouring event boundaries
        bounds.append(min(max(idx, bounds[-1]), B))
    bounds.append(B)
    return [(bounds[r], bounds[r + 1]) for r in range(world_size)]


def take_events(data: Any, lo: int, hi: int) -> Data:
    """The sub-batch holding events [lo, hi) of a collated batch (host or device tensors)."""
    batch = data.batch
    nlo = int(torch.searchsorted(batch, torch.tensor(lo, device=batch.device)))
    nhi = int(torch.searchsorted(batch, torch.tensor(hi, device=batch.device)))
    ei = data.edge_index
    keep = (ei[1] >= nlo) & (ei[1] < nhi)
    return Data(x=data.x[nlo:nhi], batch=batch[nlo:nhi] - lo, n_pulses=data.n_pulses[lo:hi],
                edge_index=ei[:, keep] - nlo)


def gather_event_outputs(local: Tensor, counts: Sequence[int]) -> Tensor:
    """all_gather of per-event outputs with different event counts per rank, in rank order."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return local
    world = dist.get_world_size()
    width = local.shape[1]
    pad = max(counts)
    buf = local.new_zeros((pad, width))
    buf[: local.shape[0]] = local
    out = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(out, buf)
    return torch.cat([o[:c] for o, c in zip(out, counts)], dim=0)


def predict_sharded(model: Callable[[Any], Tensor], data: Any, device: torch.device = None,
                    balance: str = "pulses") -> Tensor:
    """Run `model` on this rank's share of the events of `data`; return all events' outputs."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    ranges = shard_ranges(data.n_pulses, world, balance)
    lo, hi = ranges[rank]
    sub = take_events(data, lo, hi)
    if device is not None:
        sub = sub.to(device)
    out = model(sub)
    if isinstance(out, (list, tuple)):
        out = out[0]
    return gather_event_outputs(out, [h - l for l, h in ranges])
