"""Batched input pipeline: SQLite pulse tables -> packed, pinned host batches.

The reference reads ONE event per `Dataset.__getitem__` - two or more SQL queries per event
(`data/dataset/sqlite/sqlite_dataset.py:46-84` via `dataset.py:_query`), a per-event `Data` with
its own kNN graph built on the CPU (`graphs/edges/edges.py:76-84`), then
`Batch.from_data_list` in `collate_fn` (`data/dataloader.py:11-17`).  For a GPU that consumes a
1024-event batch in under a millisecond that host work is the bottleneck, so this module reads a
whole batch at once:

* one `SELECT ... WHERE event_no IN (...)` per table for all events of the batch, rows kept in
  table order inside an event (what the per-event query returns);
* the detector's column-wise standardisation applied to the packed `[N, F]` matrix (it is
  element-wise per column, so the values equal the per-event ones bit for bit);
* events with fewer than two pulses dropped, as `collate_fn` does;
* the wire format the kernels want: `x [N, F]` float32, `batch [N]` int64 (sorted), `ptr [B+1]`
  int64, `n_pulses [B]` int32, truth columns `[B]`, all in pinned host memory when CUDA is
  present, so `DevicePrefetcher` can overlap the copy with the previous batch's kernels.

No `edge_index` is produced on the host: `to_device_batch` builds the kNN graph of the whole
batch with one launch on the device (the same call as `KNNGraph.collate`).
"""

from __future__ import annotations

import ctypes as C
import os
import sqlite3
from typing import Any, Dict, Iterator, List, Optional, Sequence

import numpy as np
import torch
from torch import Tensor

from .data import Batch
from .detector import Detector


_NATIVE_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc", "libgn_sqlite.so")
_native: Optional[C.CDLL] = None


def native_reader() -> Optional[C.CDLL]:
    """The C row reader (csrc/sqlite_reader.c, built by `__graft_entry__.build()`), or None when it
    has not been built - the Python `sqlite3` module then does the same reads, ~8x slower.  This
    is host-side file I/O; the compute path has no such alternative."""
    global _native
    if _native is None and os.path.exists(_NATIVE_PATH):
        lib = C.CDLL(_NATIVE_PATH)
        lib.gn_sqlite_open.restype = C.c_void_p
        lib.gn_sqlite_open.argtypes = [C.c_char_p]
        lib.gn_sqlite_close.argtypes = [C.c_void_p]
        lib.gn_sqlite_last_error.restype = C.c_char_p
        lib.gn_sqlite_read_events.restype = C.c_int64
        lib.gn_sqlite_read_events.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int64, C.c_int,
                                              C.c_void_p, C.c_int64, C.c_void_p]
        _native = lib
    return _native


def _pinned(t: Tensor) -> Tensor:
    return t.pin_memory() if torch.cuda.is_available() else t


class PackedBatch(Batch):
    """Host-side batch in the layout the kernels consume (no edges yet)."""

    def to_device_batch(self, device: Any, nb_nearest_neighbours: int = 8,
                        columns: Sequence[int] = (0, 1, 2), non_blocking: bool = True) -> Batch:
        """Copy to `device` and build the batch's kNN graph there with one launch
        (`torch_geometric.nn.knn_graph` semantics, graphs/edges/edges.py:76-84)."""
        from . import ops
        dev = self.to(device, non_blocking=non_blocking)
        out = Batch(**{k: getattr(dev, k) for k in dev.keys()})
        out.edge_index = ops.knn_graph(out.x[:, list(columns)].contiguous(), nb_nearest_neighbours,
                                       out.batch, num_events=int(self.num_graphs))
        return out


class SQLitePulseSource:
    """Pulse maps and truth of a GraphNeT SQLite database, read a batch at a time.

    Arguments follow `graphnet.data.dataset.Dataset` (`dataset.py:213-330`): `pulsemaps` (table
    name), `features`, `truth`, `truth_table`, `index_column`, `selection` (event numbers).
    `detector` standardises the features as the reference's `GraphDefinition` does.
    """

    def __init__(self, path: str, pulsemaps: str, features: List[str], truth: Optional[List[str]] = None,
                 *, detector: Optional[Detector] = None, truth_table: str = "truth",
                 index_column: str = "event_no", selection: Optional[Sequence[int]] = None,
                 dtype: torch.dtype = torch.float32):
        assert path.endswith(".db"), f"Format of input file `{path}` is not supported."
        self._path, self._table, self._features = path, pulsemaps, list(features)
        self._truth = list(truth) if truth else []
        self._truth_table, self._index = truth_table, index_column
        self._detector, self._dtype = detector, dtype
        self._conn: Optional[sqlite3.Connection] = None
        self._handle: Optional[int] = None      # native reader's connection
        self._rows_hint = 128                   # rows per event the native buffer is sized for
        if selection is None:
            rows = self._connection().execute(
                f"SELECT {self._index} FROM {self._truth_table}").fetchall()
            selection = [int(r[0]) for r in rows]
        self._events = [int(e) for e in selection]
        # the native reader steps one `WHERE event_no = ?` statement per event: only sensible when
        # SQLite answers it from an index (GraphNeT's converters create one on the pulse tables)
        plan = self._connection().execute(
            f"EXPLAIN QUERY PLAN SELECT {self._features[0]} FROM {self._table} "
            f"WHERE {self._index} = ?", (0,)).fetchall()
        self._indexed = any("USING" in str(r[-1]) and "INDEX" in str(r[-1]) for r in plan)
        self.close()   # a forked worker must open its own connection (sqlite_dataset.py:128-140)

    # -- connection ---------------------------------------------------------------------------
    def _connection(self) -> sqlite3.Connection:
        if self._conn is None:
            self._conn = sqlite3.connect(self._path)
        return self._conn

    def close(self) -> None:
        if self._conn is not None:
            self._conn.close()
            self._conn = None
        if self._handle is not None:
            native_reader().gn_sqlite_close(self._handle)
            self._handle = None

    def __del__(self) -> None:
        try:
            self.close()
        except Exception:  # noqa: BLE001 - interpreter shutdown
            pass

    def __getstate__(self) -> Dict[str, Any]:
        state = dict(self.__dict__)
        state["_conn"] = None
        state["_handle"] = None
        return state

    def __len__(self) -> int:
        return len(self._events)

    @property
    def event_nos(self) -> List[int]:
        return list(self._events)

    # -- reading ------------------------------------------------------------------------------
    def _select(self, table: str, columns: List[str], event_nos: Sequence[int]) -> np.ndarray:
        """Rows of `table` for the events, as float64 [rows, 1 + len(columns)] (event number first),
        in table order."""
        conn = self._connection()
        parts = []
        ids = [int(e) for e in event_nos]
        for i in range(0, len(ids), 900):   # SQLite's default limit on bound variables is 999
            chunk = ids[i:i + 900]
            marks = ",".join("?" * len(chunk))
            cur = conn.execute(
                f"SELECT {self._index}, {', '.join(columns)} FROM {table} "
                f"WHERE {self._index} IN ({marks}) ORDER BY rowid", chunk)
            rows = cur.fetchall()
            if rows:
                parts.append(np.asarray(rows, dtype=np.float64))
        if not parts:
            return np.zeros((0, 1 + len(columns)), dtype=np.float64)
        return np.concatenate(parts, axis=0)

    def _read_native(self, table: str, columns: List[str], ids: np.ndarray):
        """(rows [R, len(columns)] float64 grouped by event in the order of `ids`, rows per event)
        through the C reader: one prepared `WHERE event_no = ?` statement stepped in C."""
        lib = native_reader()
        if self._handle is None:
            self._handle = lib.gn_sqlite_open(self._path.encode())
            if not self._handle:
                raise RuntimeError(lib.gn_sqlite_last_error().decode())
        sql = (f"SELECT {', '.join(columns)} FROM {table} WHERE {self._index} = ? "
               "ORDER BY rowid").encode()
        ids = np.ascontiguousarray(ids, dtype=np.int64)
        per_id = np.zeros(len(ids), dtype=np.int64)
        while True:
            cap = max(1024, self._rows_hint * len(ids))
            out = np.empty((cap, len(columns)), dtype=np.float64)
            n = lib.gn_sqlite_read_events(self._handle, sql, ids.ctypes.data, len(ids), len(columns),
                                          out.ctypes.data, cap, per_id.ctypes.data)
            if n == -2:
                self._rows_hint *= 2
                continue
            if n < 0:
                raise RuntimeError(lib.gn_sqlite_last_error().decode())
            return out[:n], per_id

    def read(self, event_nos: Sequence[int]) -> PackedBatch:
        """The events, in the order given, as one packed batch (events with < 2 pulses dropped)."""
        ids = np.asarray([int(e) for e in event_nos], dtype=np.int64)
        assert len(np.unique(ids)) == len(ids), "event numbers must be distinct within a batch"
        if self._indexed and native_reader() is not None:
            feats, counts = self._read_native(self._table, self._features, ids)
            pos = np.repeat(np.arange(len(ids)), counts)
        else:
            rows = self._select(self._table, self._features, ids)
            # position of every row's event in the requested order; stable sort keeps table order
            order = np.argsort(ids, kind="stable")
            pos = order[np.searchsorted(ids[order], rows[:, 0].astype(np.int64))]
            perm = np.argsort(pos, kind="stable")
            feats, pos = rows[perm][:, 1:], pos[perm]
            counts = np.bincount(pos, minlength=len(ids))
        keep = counts > 1                                        # data/dataloader.py:11-17
        feats = feats[keep[pos]]
        counts_kept = counts[keep]
        x = torch.tensor(feats, dtype=self._dtype)              # float64 -> dtype, as GraphDefinition
        if self._detector is not None:
            x = self._detector(x, self._features)
        B = int(keep.sum())
        ptr = torch.zeros(B + 1, dtype=torch.int64)
        ptr[1:] = torch.from_numpy(np.cumsum(counts_kept))
        batch = torch.repeat_interleave(torch.arange(B), torch.from_numpy(counts_kept))
        out = PackedBatch(x=_pinned(x.contiguous()), batch=_pinned(batch), ptr=ptr,
                          n_pulses=_pinned(torch.from_numpy(counts_kept.astype(np.int32))))
        out.event_no = torch.from_numpy(ids[keep])
        out.num_graphs = B
        if self._truth:   # B rows only: one batched query (truth tables often carry no index)
            t = self._select(self._truth_table, self._truth, ids[keep])
            torder = np.argsort(ids[keep], kind="stable")
            tpos = torder[np.searchsorted(ids[keep][torder], t[:, 0].astype(np.int64))]
            vals = np.zeros((B, len(self._truth)), dtype=np.float64)
            vals[tpos] = t[:, 1:]
            for j, name in enumerate(self._truth):
                out[name] = _pinned(torch.tensor(vals[:, j], dtype=self._dtype))
        return out

    def batches(self, batch_size: int, shuffle: bool = False, seed: int = 0,
                drop_last: bool = False, rank: int = 0, world_size: int = 1) -> Iterator[PackedBatch]:
        """Batches of `batch_size` events.  With `world_size` > 1 every rank walks its own
        contiguous share of the (optionally shuffled - same `seed` on all ranks) event list:
        events are independent, so multi-GPU inference needs nothing else (sharding.py)."""
        ev = np.asarray(self._events, dtype=np.int64)
        if shuffle:
            ev = np.random.default_rng(seed).permutation(ev)
        if world_size > 1:
            assert 0 <= rank < world_size
            lo = len(ev) * rank // world_size
            hi = len(ev) * (rank + 1) // world_size
            ev = ev[lo:hi]
        for i in range(0, len(ev), batch_size):
            chunk = ev[i:i + batch_size]
            if drop_last and len(chunk) < batch_size:
                break
            yield self.read(chunk)
