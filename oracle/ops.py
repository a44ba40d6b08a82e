"""Oracle restatements of the third-party ops on the DynEdge path (CPU, torch).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  "parity unpinned": the
libraries restated here (torch_cluster, torch_scatter, torch_geometric) are not
under /root/reference and the reference holds no known-answer test for them.

Every function cites the reference call site that reaches the library and the
SURVEY.md appendix item that states the library rule being followed.
"""

from __future__ import annotations

from typing import Callable, Optional, Tuple

import torch
from torch import Tensor


# --------------------------------------------------------------------------- #
# kNN graph
# --------------------------------------------------------------------------- #
def _pairwise_d2(c: Tensor) -> Tensor:
    """Squared L2 distances of one event, `c` = [n, D].

    Accumulated over dimensions in order, every product and every sum rounded
    separately in c.dtype (no FMA contraction): ((0 + d0^2) + d1^2) + d2^2.
    SURVEY.md A.1 (vi) / oracle rule.
    """
    n, D = c.shape
    d2 = torch.zeros((n, n), dtype=c.dtype)
    for dim in range(D):
        col = c[:, dim]
        diff = col.unsqueeze(1) - col.unsqueeze(0)  # [query, candidate]
        d2 = d2 + diff * diff
    return d2


def knn_graph(
    x: Tensor,
    k: int,
    batch: Optional[Tensor] = None,
    loop: bool = False,
) -> Tensor:
    """`torch_geometric.nn.knn_graph(x, k, batch)` (flow source_to_target).

    Reference call sites: graphs/edges/edges.py:78-82 (initial graph, called
    per event) and models/components/layers.py:70-74 (dynamic graph, batched).

    Rule: per event, per query i, candidates ordered by
    (d^2 ascending, index ascending); the first min(n, k+1) are taken when
    loop=False (k when loop=True); the pair (i, i) is dropped if present;
    output is COO int64 [2, E] with row 0 = neighbour j (source) and
    row 1 = query i (target), ordered by i then by rank.
    """
    assert x.dim() == 2
    N = x.shape[0]
    x = x.detach().cpu()
    if batch is None:
        batch = torch.zeros(N, dtype=torch.long)
    batch = batch.detach().cpu().long()
    if N == 0:
        return torch.zeros((2, 0), dtype=torch.long)
    assert bool((batch[1:] >= batch[:-1]).all()), "batch must be sorted"
    B = int(batch.max()) + 1
    ptr = torch.searchsorted(batch, torch.arange(B + 1))
    kk = k if loop else k + 1
    rows, cols = [], []
    for e in range(B):
        lo, hi = int(ptr[e]), int(ptr[e + 1])
        n = hi - lo
        if n == 0:
            continue
        d2 = _pairwise_d2(x[lo:hi])
        order = torch.sort(d2, dim=1, stable=True).indices  # ties -> low idx
        m = min(n, kk)
        nbr = order[:, :m]  # [n, m] local ids
        qry = torch.arange(n).unsqueeze(1).expand(n, m)
        if not loop:
            keep = nbr != qry
            nbr, qry = nbr[keep], qry[keep]  # row-major: by query, then rank
        else:
            nbr, qry = nbr.reshape(-1), qry.reshape(-1)
        rows.append(nbr + lo)
        cols.append(qry + lo)
    if not rows:
        return torch.zeros((2, 0), dtype=torch.long)
    return torch.stack([torch.cat(rows), torch.cat(cols)], dim=0)


def knn_graph_padded(x: Tensor, k: int, batch: Tensor) -> Tensor:
    """Vectorised twin of `knn_graph` (same result, used for CPU timing).

    Events are padded to a dense [B, nmax, nmax] distance tensor; ordering by
    (d^2, index) is obtained from one integer sort of a composite key (the
    IEEE-754 bit pattern of a non-negative float is monotone in its value).
    """
    x = x.detach().cpu().float()
    batch = batch.detach().cpu().long()
    N = x.shape[0]
    B = int(batch.max()) + 1
    counts = torch.bincount(batch, minlength=B)
    ptr = torch.zeros(B + 1, dtype=torch.long)
    ptr[1:] = counts.cumsum(0)
    nmax = int(counts.max())
    local = torch.arange(N) - ptr[batch]
    D = x.shape[1]
    pad = torch.zeros((B, nmax, D), dtype=x.dtype)
    pad[batch, local] = x
    valid = torch.zeros((B, nmax), dtype=torch.bool)
    valid[batch, local] = True
    d2 = torch.zeros((B, nmax, nmax), dtype=x.dtype)
    for dim in range(D):
        col = pad[:, :, dim]
        diff = col.unsqueeze(2) - col.unsqueeze(1)
        d2 = d2 + diff * diff
    key = d2.view(torch.int32).long() << 32
    key = key + torch.arange(nmax).view(1, 1, nmax)
    big = torch.iinfo(torch.int64).max
    key = torch.where(valid.unsqueeze(1), key, torch.full_like(key, big))
    m = min(nmax, k + 1)
    top = torch.topk(key, m, dim=2, largest=False, sorted=True).values
    nbr = top & 0xFFFFFFFF  # [B, nmax, m] local neighbour ids
    rank = torch.arange(m).view(1, 1, m)
    ok = (
        valid.unsqueeze(2)
        & (rank < counts.view(B, 1, 1).clamp(max=k + 1))
        & (nbr != torch.arange(nmax).view(1, nmax, 1))
    )
    qry = torch.arange(nmax).view(1, nmax, 1).expand(B, nmax, m)
    off = ptr[:-1].view(B, 1, 1)
    row = (nbr + off)[ok]
    col = (qry + off)[ok]
    return torch.stack([row, col], dim=0)


# --------------------------------------------------------------------------- #
# EdgeConv
# --------------------------------------------------------------------------- #
def edge_conv(
    x: Tensor,
    edge_index: Tensor,
    nn: Callable[[Tensor], Tensor],
    aggr: str = "add",
) -> Tensor:
    """`torch_geometric.nn.EdgeConv(nn, aggr).forward(x, edge_index)`.

    Reference call site: models/components/layers.py:67 (DynEdgeConv.forward,
    `super().forward(x, edge_index)`); DynEdge builds it with aggr="add"
    (models/gnn/dynedge.py:206-211).

    Rule: x_j = x[edge_index[0]], x_i = x[edge_index[1]],
    message = nn(cat([x_i, x_j - x_i], -1)), aggregated per target
    edge_index[1] over N rows; rows without edges are 0.
    """
    src, dst = edge_index[0], edge_index[1]
    N = x.shape[0]
    if src.numel() == 0:
        probe = nn(torch.zeros((1, 2 * x.shape[1]), dtype=x.dtype))
        return torch.zeros((N, probe.shape[1]), dtype=x.dtype)
    x_i = x.index_select(0, dst)
    x_j = x.index_select(0, src)
    msg = nn(torch.cat([x_i, x_j - x_i], dim=-1))
    if aggr in ("add", "sum"):
        return scatter_sum(msg, dst, dim_size=N)
    if aggr == "mean":
        return scatter_mean(msg, dst, dim_size=N)
    if aggr == "max":
        return scatter_max(msg, dst, dim_size=N)[0]
    if aggr == "min":
        return scatter_min(msg, dst, dim_size=N)[0]
    raise ValueError(aggr)


# --------------------------------------------------------------------------- #
# torch_scatter
# --------------------------------------------------------------------------- #
def _dim_size(index: Tensor, dim_size: Optional[int]) -> int:
    if dim_size is not None:
        return dim_size
    return int(index.max()) + 1 if index.numel() else 0


def scatter_sum(
    src: Tensor, index: Tensor, dim_size: Optional[int] = None
) -> Tensor:
    """`torch_scatter.scatter_sum(src, index, dim=0)` (dynedge.py:14-19,258).

    zeros + index_add in element order.
    """
    M = _dim_size(index, dim_size)
    out = torch.zeros((M,) + tuple(src.shape[1:]), dtype=src.dtype)
    return out.index_add_(0, index, src)


def scatter_mean(
    src: Tensor, index: Tensor, dim_size: Optional[int] = None
) -> Tensor:
    """`torch_scatter.scatter_mean` (dynedge.py:279,258): sum / max(count,1)."""
    M = _dim_size(index, dim_size)
    s = scatter_sum(src, index, M)
    cnt = torch.zeros(M, dtype=src.dtype).index_add_(
        0, index, torch.ones(index.shape[0], dtype=src.dtype)
    )
    cnt = cnt.clamp(min=1)
    return s / cnt.view((M,) + (1,) * (src.dim() - 1))


def _scatter_extreme(
    src: Tensor, index: Tensor, dim_size: Optional[int], mode: str
) -> Tuple[Tensor, Tensor]:
    M = _dim_size(index, dim_size)
    shape = (M,) + tuple(src.shape[1:])
    out = torch.zeros(shape, dtype=src.dtype)
    idx = index.view((-1,) + (1,) * (src.dim() - 1)).expand_as(src)
    out.scatter_reduce_(0, idx, src, reduce=mode, include_self=False)
    # arg = smallest element index attaining the extreme; N for empty segments
    N = src.shape[0]
    hit = src == out.index_select(0, index)
    pos = torch.arange(N).view((-1,) + (1,) * (src.dim() - 1)).expand_as(src)
    pos = torch.where(hit, pos, torch.full_like(pos, N))
    arg = torch.full(shape, N, dtype=torch.long)
    arg.scatter_reduce_(0, idx, pos, reduce="amin", include_self=True)
    return out, arg


def scatter_min(
    src: Tensor, index: Tensor, dim_size: Optional[int] = None
) -> Tuple[Tensor, Tensor]:
    """`torch_scatter.scatter_min` -> (values, arg); empty segments -> 0."""
    return _scatter_extreme(src, index, dim_size, "amin")


def scatter_max(
    src: Tensor, index: Tensor, dim_size: Optional[int] = None
) -> Tuple[Tensor, Tensor]:
    """`torch_scatter.scatter_max` -> (values, arg); empty segments -> 0."""
    return _scatter_extreme(src, index, dim_size, "amax")


# --------------------------------------------------------------------------- #
# homophily
# --------------------------------------------------------------------------- #
def homophily(edge_index: Tensor, y: Tensor, batch: Tensor) -> Tensor:
    """`torch_geometric.utils.homophily(edge_index, y, batch)`, method "edge".

    Reference call site: models/utils.py:29-32.  Rule: per
    edge 1.0 if y[row] == y[col] (exact equality) else 0.0, averaged per event
    of the target node; events without edges give 0.
    """
    row, col = edge_index[0], edge_index[1]
    out = torch.zeros(row.shape[0], dtype=torch.float32)
    out[y[row] == y[col]] = 1.0
    M = int(batch.max()) + 1
    return scatter_mean(out, batch[col], dim_size=M)
