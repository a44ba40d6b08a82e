"""GPU: SQLite -> packed pinned batch -> device kNN -> DynEdge (SURVEY.md 8(f).4 + 8(f).1)."""
import sqlite3

import pytest
import torch

import oracle
from conftest import rel_l2
from test_io import _make_db

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_database_to_prediction(tmp_path):
    import graphnet_b200 as gb
    from graphnet_b200.data import DevicePrefetcher
    from graphnet_b200.detector import FEATURES_ICECUBE86, IceCube86
    from graphnet_b200.io import SQLitePulseSource
    path = str(tmp_path / "events.db")
    _make_db(path, [30, 2, 45, 1, 17, 60, 9, 33], seed=3)
    con = sqlite3.connect(path)
    con.execute("CREATE INDEX pulses_event_no ON pulses(event_no)")
    con.commit()
    con.close()
    src = SQLitePulseSource(path, "pulses", FEATURES_ICECUBE86, ["energy"], detector=IceCube86())
    pools = ["min", "max", "mean", "sum"]
    spec = oracle.DynEdgeSpec(nb_inputs=7, global_pooling_schemes=pools)
    sd = oracle.init_dynedge_state_dict(spec, seed=0)
    model = gb.DynEdge(7, global_pooling_schemes=pools).to(DEV)
    model.eval()   # inference: the fp32 / fp16 paths have no backward
    model.load_state_dict(sd)
    model.train(False)
    seen = 0
    for host in src.batches(4):
        assert host.x.is_pinned() and host.batch.is_pinned()
        dev = host.to_device_batch(DEV, nb_nearest_neighbours=8)
        want_ei = oracle.knn_graph(host.x[:, :3], 8, host.batch)
        assert torch.equal(dev.edge_index.cpu(), want_ei)
        with torch.no_grad():
            got = model(dev).cpu()
        want = oracle.dynedge_forward(spec, sd, host.x, want_ei, host.batch, host.n_pulses)
        assert rel_l2(got, want) < 1e-4
        assert dev.energy.shape == (host.num_graphs,)
        seen += host.num_graphs
    assert seen == 7   # the single-pulse event is dropped (collate_fn)
    # the same batches through the prefetcher (copy of batch i+1 under batch i's kernels)
    n = sum(int(b.x.shape[0]) for b in DevicePrefetcher(src.batches(4), DEV))
    assert n == sum(c for c in [30, 2, 45, 17, 60, 9, 33])
