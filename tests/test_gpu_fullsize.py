"""BASELINE.json's full batch sizes, checked through size-independent properties (the CPU oracle
needs minutes there): structure of the kNN graph, and invariance of the forward under splitting
the batch into sub-batches (events are independent - dynedge.py never mixes `batch` ids)."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")


def _batch(events, pulses, detector="icecube86", seed=5):
    import graphnet_b200 as gb
    from graphnet_b200.synthetic import make_pulse_batch
    pb = make_pulse_batch(events, pulses, detector=detector, seed=seed)
    return pb, pb.x.to(DEV), pb.batch.to(DEV), pb.n_pulses.to(DEV)


def _check_graph(ei, x3, batch, k):
    src, dst = ei[0], ei[1]
    N = x3.shape[0]
    assert int(ei.min()) >= 0 and int(ei.max()) < N
    assert bool((dst[1:] >= dst[:-1]).all()), "targets must be grouped in node order"
    assert bool((src != dst).all()), "self loops are dropped"
    assert bool((batch[src] == batch[dst]).all()), "edges never cross events"
    deg = torch.bincount(dst, minlength=N)
    n_ev = torch.bincount(batch)[batch]
    # k+1 nearest including self, self dropped if present: min(n, k+1) - 1 <= deg <= min(n, k+1)
    hi = torch.minimum(n_ev, torch.full_like(n_ev, k + 1))
    assert bool((deg <= hi).all()) and bool((deg >= hi - 1).all())
    # per target: (d2, index) strictly increasing along its edge list
    d = x3[src] - x3[dst]
    d2 = ((0 + d[:, 0] * d[:, 0]) + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]
    same = dst[1:] == dst[:-1]
    ok = (d2[1:] > d2[:-1]) | ((d2[1:] == d2[:-1]) & (src[1:] > src[:-1]))
    assert bool((ok | ~same).all()), "neighbours must be ordered by (distance, index)"
    # nothing closer was left out: the worst kept neighbour is no farther than any non-neighbour
    # of the same event (checked on a sample of targets)
    g = torch.Generator().manual_seed(0)
    for t in torch.randint(0, N, (64,), generator=g).tolist():
        members = (batch == batch[t]).nonzero().flatten()
        dd = x3[members] - x3[t]
        dist = ((0 + dd[:, 0] * dd[:, 0]) + dd[:, 1] * dd[:, 1]) + dd[:, 2] * dd[:, 2]
        mine = src[dst == t]
        kept = set(mine.tolist()) | {t}
        order = sorted(range(len(members)), key=lambda i: (float(dist[i]), int(members[i])))
        want = {int(members[i]) for i in order[:k + 1]}
        assert want - {t} == kept - {t} or want == kept, f"target {t}"


@pytest.mark.parametrize("events,pulses,detector,k", [(1024, 60.0, "icecube86", 8),
                                                       (256, 500.0, "upgrade", 16)])
def test_knn_graph_structure_full_size(events, pulses, detector, k):
    import graphnet_b200 as gb
    pb, x, batch, _ = _batch(events, pulses, detector)
    x3 = x[:, :3].contiguous()
    ei = gb.knn_graph(x3, k, batch, num_events=events)
    _check_graph(ei, x3, batch, k)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("events,pulses,detector,k,F,pools", [
    (1024, 60.0, "icecube86", 8, 7, ["min", "max", "mean", "sum"]),
    (192, 500.0, "upgrade", 16, 14, ["min", "max", "mean"])])
def test_forward_is_invariant_under_batch_splitting(precision, events, pulses, detector, k, F, pools):
    import graphnet_b200 as gb
    from graphnet_b200.sharding import take_events
    pb, x, batch, n_pulses = _batch(events, pulses, detector)
    ei = gb.knn_graph(x[:, :3].contiguous(), k, batch, num_events=events)
    data = gb.Data(x=x, edge_index=ei, batch=batch, n_pulses=n_pulses)
    torch.manual_seed(2)
    m = gb.DynEdge(F, nb_neighbours=k, global_pooling_schemes=pools, precision=precision).to(DEV)
    m.train(False)
    with torch.no_grad():
        whole = m(data)
        cuts = [0, events // 3, events // 2 + 1, events]
        parts = [m(take_events(data, lo, hi)) for lo, hi in zip(cuts, cuts[1:])]
    got = torch.cat(parts, 0)
    assert whole.shape == (events, 128) and bool(torch.isfinite(whole).all())
    # every kernel treats events (and rows) independently and sums in a fixed order per row, so
    # the split must reproduce the whole batch bit for bit
    assert torch.equal(got, whole), float((got - whole).abs().max())
