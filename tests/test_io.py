"""CPU: the batched SQLite reader (graphnet_b200/io.py) against per-event reads."""
import os
import sqlite3

import numpy as np
import pytest
import torch

from conftest import load_golden
from graphnet_b200.detector import FEATURES_ICECUBE86, IceCube86
from graphnet_b200.io import SQLitePulseSource

REF_DB = ("/root/reference/data/tests/sqlite/oscNext_genie_level7_v02/"
          "oscNext_genie_level7_v02_first_5_frames.db")


def _make_db(path, counts, seed=0):
    rng = np.random.default_rng(seed)
    con = sqlite3.connect(path)
    cols = ", ".join(f"{f} REAL" for f in FEATURES_ICECUBE86)
    con.execute(f"CREATE TABLE pulses (event_no INTEGER, {cols})")
    con.execute("CREATE TABLE truth (event_no INTEGER, energy REAL, zenith REAL)")
    rows, truth, per_event = [], [], {}
    ev_ids = [100 + 7 * i for i in range(len(counts))]
    # interleave the events' rows in the table: a reader must group them itself
    pending = [(ev, n) for ev, n in zip(ev_ids, counts)]
    while any(n > 0 for _, n in pending):
        for i, (ev, n) in enumerate(pending):
            if n == 0:
                continue
            r = [float(rng.normal(0, 300)), float(rng.normal(0, 300)), float(rng.normal(-200, 200)),
                 float(rng.uniform(9000, 14000)), float(rng.lognormal(0, 1)),
                 float(rng.choice([1.0, 1.35])), 0.0444]
            rows.append((ev, *r))
            per_event.setdefault(ev, []).append(r)
            pending[i] = (ev, n - 1)
    for ev in ev_ids:
        truth.append((ev, float(rng.uniform(1, 1e4)), float(rng.uniform(0, np.pi))))
    con.executemany(f"INSERT INTO pulses VALUES ({','.join('?' * 8)})", rows)
    con.executemany("INSERT INTO truth VALUES (?,?,?)", truth)
    con.commit()
    con.close()
    return ev_ids, per_event, {t[0]: t[1:] for t in truth}


@pytest.mark.parametrize("indexed", [False, True])
def test_batched_read_equals_per_event_reads(tmp_path, indexed):
    """indexed=True: the C reader (one prepared statement per batch, csrc/sqlite_reader.c) when it
    is built; otherwise - and for a table without an index - Python's sqlite3 with one IN query."""
    path = str(tmp_path / "events.db")
    counts = [5, 1, 12, 0, 2, 40, 3]
    ev_ids, per_event, truth = _make_db(path, counts)
    if indexed:
        con = sqlite3.connect(path)
        con.execute("CREATE INDEX pulses_event_no ON pulses(event_no)")
        con.commit()
        con.close()
    src = SQLitePulseSource(path, "pulses", FEATURES_ICECUBE86, ["energy", "zenith"],
                            detector=IceCube86())
    assert src._indexed == indexed
    assert src.event_nos == ev_ids and len(src) == 7
    want_order = [ev_ids[5], ev_ids[0], ev_ids[1], ev_ids[3], ev_ids[2], ev_ids[6]]
    b = src.read(want_order)
    kept = [ev for ev in want_order if len(per_event.get(ev, [])) > 1]   # collate_fn drops n <= 1
    assert b.event_no.tolist() == kept and b.num_graphs == len(kept)
    det = IceCube86()
    xs = [det(torch.tensor(np.asarray(per_event[ev], dtype=np.float64), dtype=torch.float32),
              FEATURES_ICECUBE86) for ev in kept]
    assert torch.equal(b.x, torch.cat(xs))
    assert b.x.dtype == torch.float32 and b.n_pulses.dtype == torch.int32
    assert b.n_pulses.tolist() == [len(per_event[ev]) for ev in kept]
    assert b.ptr.tolist() == np.concatenate([[0], np.cumsum(b.n_pulses.numpy())]).tolist()
    assert torch.equal(b.batch, torch.repeat_interleave(torch.arange(len(kept)), b.n_pulses.long()))
    assert torch.allclose(b.energy, torch.tensor([truth[ev][0] for ev in kept], dtype=torch.float32))
    assert torch.allclose(b.zenith, torch.tensor([truth[ev][1] for ev in kept], dtype=torch.float32))
    # batches(): every event once, batch boundaries respected
    seen = [e for bb in src.batches(3) for e in bb.event_no.tolist()]
    assert seen == [ev for ev in ev_ids if len(per_event.get(ev, [])) > 1]
    shuffled = [e for bb in src.batches(3, shuffle=True, seed=1) for e in bb.event_no.tolist()]
    assert sorted(shuffled) == sorted(seen) and shuffled != seen


def test_ranks_read_disjoint_shares(tmp_path):
    path = str(tmp_path / "events.db")
    ev_ids, per_event, _ = _make_db(path, [4, 6, 3, 9, 2, 5, 7, 8, 3, 4, 6])
    src = SQLitePulseSource(path, "pulses", FEATURES_ICECUBE86)
    for shuffle in (False, True):
        shares = [[e for b in src.batches(2, shuffle=shuffle, seed=5, rank=r, world_size=3)
                   for e in b.event_no.tolist()] for r in range(3)]
        assert sorted(sum(shares, [])) == sorted(ev_ids)
        assert max(len(s) for s in shares) - min(len(s) for s in shares) <= 1
    one = [e for b in src.batches(2) for e in b.event_no.tolist()]
    assert one == ev_ids


def test_source_survives_pickling_without_connection(tmp_path):
    import pickle
    path = str(tmp_path / "events.db")
    ev_ids, _, _ = _make_db(path, [3, 4])
    src = SQLitePulseSource(path, "pulses", FEATURES_ICECUBE86)
    src.read(ev_ids)                   # opens a connection
    clone = pickle.loads(pickle.dumps(src))
    assert clone._conn is None and clone.read(ev_ids).num_graphs == 2


@pytest.mark.skipif(not os.path.exists(REF_DB), reason="the reference's bundled database is not here")
def test_reference_database_gives_the_golden_batch():
    """The reference's own oscNext test database through the batched reader = the inputs the golden
    forward was recorded on (tests/golden/make_golden.py reads it event by event)."""
    g = load_golden("oscnext5")
    src = SQLitePulseSource(REF_DB, "SRTInIcePulses", FEATURES_ICECUBE86, detector=IceCube86())
    b = src.read(sorted(src.event_nos))
    assert torch.equal(b.x, g["x"]) and torch.equal(b.batch, g["batch"])
    assert torch.equal(b.n_pulses, g["n_pulses"].to(torch.int32))
