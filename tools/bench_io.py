"""Host throughput of the batched SQLite reader vs one query per event (the reference's access
pattern, sqlite_dataset.py:46-84).  CPU only.   python tools/bench_io.py [events]"""
import os
import sqlite3
import sys
import tempfile
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graphnet_b200.detector import FEATURES_ICECUBE86, IceCube86  # noqa: E402
from graphnet_b200.io import SQLitePulseSource  # noqa: E402

n_events = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
rng = np.random.default_rng(0)
counts = np.maximum(2, rng.poisson(60, n_events))
N = int(counts.sum())
ev = np.repeat(np.arange(n_events), counts)
feat = rng.normal(size=(N, 7))
feat[:, 4] = np.abs(feat[:, 4]) + 0.1
path = os.path.join(tempfile.mkdtemp(), "bench.db")
con = sqlite3.connect(path)
con.execute("CREATE TABLE pulses (event_no INTEGER, " + ", ".join(f"{f} REAL" for f in FEATURES_ICECUBE86) + ")")
con.execute("CREATE TABLE truth (event_no INTEGER, energy REAL)")
con.executemany("INSERT INTO pulses VALUES (?,?,?,?,?,?,?,?)",
                [(int(e), *map(float, r)) for e, r in zip(ev, feat)])
con.executemany("INSERT INTO truth VALUES (?,?)", [(i, 1.0) for i in range(n_events)])
con.execute("CREATE INDEX idx ON pulses(event_no)")
con.commit()
con.close()

src = SQLitePulseSource(path, "pulses", FEATURES_ICECUBE86, ["energy"], detector=IceCube86())
t0 = time.perf_counter()
nb = sum(b.num_graphs for b in src.batches(1024))
t1 = time.perf_counter()
print(f"batched reader: {nb / (t1 - t0):,.0f} events/s ({N / (t1 - t0) / 1e6:.2f} M pulses/s), one host thread")

det = IceCube86()
con = sqlite3.connect(path)
sample = min(n_events, 2048)
t0 = time.perf_counter()
for e in range(sample):
    rows = con.execute(f"SELECT {', '.join(FEATURES_ICECUBE86)} FROM pulses WHERE event_no = {e}").fetchall()
    con.execute(f"SELECT energy FROM truth WHERE event_no = {e}").fetchall()
    det(torch.tensor(np.asarray(rows), dtype=torch.float32), FEATURES_ICECUBE86)
t1 = time.perf_counter()
print(f"one query per event (reference access pattern, no graph building): {sample / (t1 - t0):,.0f} events/s")
