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
    m.eval()   # inference: the fp32 / fp16 paths have no backward
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


# --------------------------------------------------------------------------- #
# BASELINE configs 2 and 4 at the BENCHMARKED batch size against the committed golden outputs
# (tests/golden/make_golden.py --fullsize-only: the reference's own dynedge.py on CPU; inputs are
# regenerated from the seed, the golden holds outputs and per-event digests of the dynamic graphs)
# --------------------------------------------------------------------------- #
def _graph_digest(ei, batch, ptr, num_events):
    """tests/golden/make_golden.py:graph_digest on the device (int64 arithmetic wraps alike)."""
    src, dst = ei[0], ei[1]
    ev = batch[dst]
    key = (src - ptr[ev]) * 4096 + (dst - ptr[ev]) + 1
    h = key * 0x9E3779B97F4A7C1 + (key * key) * 0x85EBCA6B
    out = torch.zeros(num_events, dtype=torch.int64, device=ei.device)
    out.index_add_(0, ev, h)
    return out


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("name", ["c2_1024_out", "c4_256_out"])
def test_full_size_forward_against_golden(name, precision):
    import graphnet_b200 as gb
    import graphnet_b200.ops as ops
    import oracle
    from conftest import load_golden, rel_l2
    from graphnet_b200.dynedge_bf16 import forward_bf16
    from graphnet_b200.synthetic import make_pulse_batch
    g = load_golden(name)
    events, k = int(g["events"]), int(g["k"])
    pb = make_pulse_batch(events, float(g["pulses"]), detector=g["detector"], seed=int(g["data_seed"]))
    x, batch, n_pulses = pb.x.to(DEV), pb.batch.to(DEV), pb.n_pulses.to(DEV)
    ei = gb.knn_graph(x[:, :3].contiguous(), k, batch, num_events=events)
    # the regenerated inputs are the generator's (same numpy stream, same initial graph)
    dig = g["input_digest"]
    assert float(pb.x.double().sum()) == float(dig[0]) and pb.x.shape[0] == int(dig[1])
    assert ei.shape[1] == int(dig[2]) and float(ei.double().sum()) == float(dig[3])
    spec = oracle.DynEdgeSpec(nb_inputs=x.shape[1], **g["dyn_kwargs"])
    sd = oracle.init_dynedge_state_dict(spec, seed=int(g["seed"]))
    m = gb.DynEdge(x.shape[1], precision=precision, **g["dyn_kwargs"])
    m.load_state_dict(sd)
    m = m.to(DEV).eval()
    data = gb.Data(x=x, edge_index=ei, batch=batch, n_pulses=n_pulses)
    with torch.no_grad():
        if precision == "fp32":
            out, inter = m._forward_f32(data, return_intermediates=True)
        else:
            out, inter = forward_bf16(m, data, return_intermediates=True)
    want = g["out"].to(DEV)
    assert out.shape == want.shape and bool(torch.isfinite(out).all())
    if precision == "bf16":
        # north_star: 2e-2 relative for the 16-bit path, at the benchmarked size
        assert rel_l2(out.cpu(), want.cpu()) < 2e-2
        return
    # fp32: the dynamic graphs come from latents that differ from the CPU run by summation order,
    # so a few near-tied neighbours swap; events whose three dynamic graphs equal the golden run's
    # must agree to 1e-4, all events together more loosely
    ptr = pb.ptr.to(DEV)
    same = torch.ones(events, dtype=torch.bool, device=DEV)
    for i in range(3):
        mine = _graph_digest(ops.ell_to_coo(inter["graphs"][i]), batch, ptr, events)
        same &= mine == g["graph_digest"][i].to(DEV)
    frac = float(same.float().mean())
    assert frac > 0.5, f"only {frac:.2%} of the events kept the golden dynamic graphs"
    assert rel_l2(out[same].cpu(), want[same].cpu()) < 1e-4, f"swap-free events ({frac:.1%})"
    assert rel_l2(out.cpu(), want.cpu()) < 5e-3
