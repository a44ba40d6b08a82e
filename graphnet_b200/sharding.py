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


def take_events(data: Any, lo: int, hi: int, rebase: bool = True) -> Data:
    """The sub-batch holding events [lo, hi) of a collated batch (host or device tensors).
    `rebase=False` leaves `batch` (and `edge_index`) as views with the parent's numbering - the
    caller subtracts `lo` (`sub.node_offset` for edges) after the copy to the device, so that
    pinned host tensors stay pinned views and the copy stays asynchronous."""
    batch = data.batch
    nlo = int(torch.searchsorted(batch, torch.tensor(lo, device=batch.device)))
    nhi = int(torch.searchsorted(batch, torch.tensor(hi, device=batch.device)))
    b = batch[nlo:nhi]
    sub = Data(x=data.x[nlo:nhi], batch=b - lo if rebase else b, n_pulses=data.n_pulses[lo:hi])
    ei = getattr(data, "edge_index", None)
    if ei is not None:   # a batch without edges gets its graph on the device (DynEdge._graph_inputs)
        keep = (ei[1] >= nlo) & (ei[1] < nhi)
        sub.edge_index = ei[:, keep] - nlo
    return sub


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
    if device is not None and getattr(data, "edge_index", None) is None:
        # views of the (pinned) host tensors, copied asynchronously, re-numbered on the device
        sub = take_events(data, lo, hi, rebase=False).to(device, non_blocking=True)
        sub.batch = sub.batch - lo
    else:
        sub = take_events(data, lo, hi)
        if device is not None:
            sub = sub.to(device)
    out = model(sub)
    if isinstance(out, (list, tuple)):
        out = out[0]
    return gather_event_outputs(out, [h - l for l, h in ranges])


def allreduce_gradients(params: Sequence[torch.nn.Parameter], average: bool = True) -> int:
    """Data-parallel training: one all-reduce over every gradient of the replica.

    The reference trains under Lightning's DDP strategy (easy_model.py:101-124), i.e. an NCCL
    all-reduce of the gradients per step; DynEdge has 1.38 M parameters (5.5 MB fp32), so the
    whole set goes out as ONE flat buffer after the backward pass - launch latency, not
    bandwidth, is what this exchange costs on NVLink.  Parameters without a gradient count as
    zeros (every rank must contribute the same layout).  Returns the bytes reduced.
    """
    params = [p for p in params if p.requires_grad]
    if not params or not (dist.is_available() and dist.is_initialized()) or \
            dist.get_world_size() == 1:
        return 0
    grads = []
    for p in params:
        if p.grad is None:            # every rank must contribute the same layout
            p.grad = torch.zeros_like(p)
        grads.append(p.grad)
    # three launches whatever the number of parameters: one flatten, one all-reduce (NCCL averages
    # in the reduction itself), one multi-tensor copy back into the .grad tensors.  (A copy_ per
    # parameter cost ~46 launches on a step that is bound by the host's enqueue rate: 0.94
    # scaling efficiency at 8 GPUs.)
    flat = torch.cat([g.reshape(-1).float() for g in grads])
    op = dist.ReduceOp.AVG if average and dist.get_backend() == "nccl" else dist.ReduceOp.SUM
    dist.all_reduce(flat, op=op)
    if average and op == dist.ReduceOp.SUM:
        flat /= dist.get_world_size()
    views, off = [], 0
    for g in grads:
        n = g.numel()
        views.append(flat[off:off + n].view_as(g))
        off += n
    if all(g.dtype == torch.float32 for g in grads):
        torch._foreach_copy_(grads, views)
    else:
        for g, v in zip(grads, views):
            g.copy_(v)
    return flat.numel() * 4
