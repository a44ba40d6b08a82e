"""GPU: segmented kNN graph construction through the C ABI - bit-exact vs the oracle."""
import pytest
import torch

import oracle
from conftest import load_golden, random_events, GOLDEN_CASES

pytestmark = pytest.mark.gpu


def dev():
    return torch.device("cuda:0")


def gpu_knn(x, k, batch, loop=False):
    import graphnet_b200 as g
    return g.knn_graph(x.to(dev()), k, None if batch is None else batch.to(dev()), loop=loop).cpu()


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_golden_initial_graphs(name):
    g = load_golden(name)
    ei = gpu_knn(g["x"][:, :3].contiguous(), int(g["k"]), g["batch"])
    assert ei.dtype == torch.int64
    assert torch.equal(ei, g["edge_index"])


@pytest.mark.parametrize("name", ["upgrade5", "prometheus50", "adversarial"])
def test_golden_dynamic_graphs_from_golden_latents(name):
    """kNN on the first 3 latent channels of each recorded conv output (layers.py:70-74)."""
    g = load_golden(name)
    for i in range(4):
        x = g[f"conv{i}"][:, :3].contiguous()
        ei = gpu_knn(x, 8, g["batch"])
        assert torch.equal(ei, g[f"graph{i + 1}"]), f"layer {i}"


@pytest.mark.parametrize("k", [1, 2, 8, 16, 32, 40])
@pytest.mark.parametrize("dims", [1, 2, 3, 4, 7])
def test_random_events_k_dims(k, dims):
    x, batch, _ = random_events(24, 1, 90, F=7, seed=k * 10 + dims)
    xs = x[:, :dims].contiguous()
    assert torch.equal(gpu_knn(xs, k, batch), oracle.knn_graph(xs, k, batch))


def test_loop_true_and_no_batch():
    x, batch, _ = random_events(6, 3, 40, seed=3)
    xs = x[:, :3].contiguous()
    assert torch.equal(gpu_knn(xs, 4, batch, loop=True), oracle.knn_graph(xs, 4, batch, loop=True))
    one = xs[:37].contiguous()
    assert torch.equal(gpu_knn(one, 8, None), oracle.knn_graph(one, 8, None))


def test_degenerate_events():
    # single-pulse events, identical points, exact ties on a lattice, n == k+1
    parts = [torch.zeros(1, 3), torch.full((13, 3), 0.5), torch.randn(9, 3),
             torch.round(torch.randn(50, 3) * 2) / 2, torch.zeros(1, 3), torch.randn(2, 3)]
    x = torch.cat(parts)
    counts = torch.tensor([p.shape[0] for p in parts])
    batch = torch.repeat_interleave(torch.arange(len(parts)), counts)
    want = oracle.knn_graph(x, 8, batch)
    got = gpu_knn(x, 8, batch)
    assert torch.equal(got, want)
    deg = torch.bincount(got[1], minlength=x.shape[0])
    assert deg[0] == 0 and deg[1:14].max() == 9  # variable degree 0..k+1


def test_large_events_cross_chunk():
    """Events larger than the shared-memory staging chunk (2048 points at 3 dims = 2730)."""
    g = torch.Generator().manual_seed(1)
    counts = torch.tensor([3000, 5, 2731, 2730])
    x = torch.randn(int(counts.sum()), 3, generator=g)
    x[100:140] = x[100]
    batch = torch.repeat_interleave(torch.arange(4), counts)
    assert torch.equal(gpu_knn(x, 8, batch), oracle.knn_graph(x, 8, batch))


@pytest.mark.parametrize("k", [8, 16])
@pytest.mark.parametrize("counts", [
    [500, 480, 700, 130, 70, 20],          # mean > 128: several CTAs per event, queued scan
    [65, 100, 159, 160, 161, 64, 63, 2],   # around the switch points of the three scan shapes
    [900] + [30] * 40,                     # one large event in a batch of small ones (one CTA)
    [2048, 1500],                          # several staging chunks and several CTAs per event
])
def test_scan_shapes_with_ties(counts, k):
    """Two threads per query (n <= 64), immediate insertion, and the queued scan of large
    events (csrc/knn.cu) must all give the (distance, index) order; lattice coordinates and
    copied rows put many exact ties across the k-th place."""
    g = torch.Generator().manual_seed(sum(counts) + k)
    counts_t = torch.tensor(counts)
    n_all = int(counts_t.sum())
    x = torch.round(torch.randn(n_all, 3, generator=g) * 3) / 2       # coarse lattice: ties
    x[n_all // 3: n_all // 3 + 40] = x[n_all // 3]                     # 40 coincident pulses
    fine = torch.rand(n_all, generator=g) < 0.5
    x[fine] += torch.randn(int(fine.sum()), 3, generator=g) * 1e-3
    batch = torch.repeat_interleave(torch.arange(len(counts)), counts_t)
    assert torch.equal(gpu_knn(x.contiguous(), k, batch), oracle.knn_graph(x, k, batch))


@pytest.mark.parametrize("k", [2, 8])
@pytest.mark.parametrize("lo,hi,outliers", [(1, 12, []), (2, 30, [70, 200]), (16, 16, []), (1, 3, [64, 65])])
def test_small_events_share_a_cta(lo, hi, outliers, k):
    """Mean event size below 32: several consecutive events are staged and searched by one CTA
    (csrc/knn.cu, p.group > 1); an event above 64 pulses inside such a batch sends its group
    down the one-event-at-a-time path.  Same (distance, index) order either way."""
    g = torch.Generator().manual_seed(lo * 100 + hi + k)
    counts = torch.randint(lo, hi + 1, (301,), generator=g)
    for i, n in enumerate(outliers):
        counts[37 + 50 * i] = n
    n_all = int(counts.sum())
    x = torch.round(torch.randn(n_all, 3, generator=g) * 2) / 2       # lattice: exact ties
    fine = torch.rand(n_all, generator=g) < 0.5
    x[fine] += torch.randn(int(fine.sum()), 3, generator=g) * 1e-3
    batch = torch.repeat_interleave(torch.arange(len(counts)), counts)
    assert n_all / len(counts) < 32
    assert torch.equal(gpu_knn(x.contiguous(), k, batch), oracle.knn_graph(x, k, batch))


def test_strided_latent_columns_via_ell():
    """The in-model call: columns of a wide row-major matrix, ELL output, no COO."""
    import graphnet_b200.ops as ops
    x, batch, _ = random_events(10, 2, 60, F=7, seed=8)
    wide = torch.randn(x.shape[0], 300)
    wide[:, 40:43] = x[:, :3]
    ptr, flags = ops.batch_to_ptr(batch.to(dev()), 10)
    assert int(flags.item()) == 0
    gph = ops.knn_graph_ell(wide.to(dev()), 8, ptr, [40, 41, 42])
    ei = ops.ell_to_coo(gph).cpu()
    assert torch.equal(ei, oracle.knn_graph(x[:, :3], 8, batch))
    assert int(gph.deg.max()) <= 9 and gph.col.shape == (x.shape[0], 9)


def test_full_size_properties():
    """C2-sized batch: degree rules, sortedness, symmetry-free invariants (no oracle)."""
    from graphnet_b200.synthetic import make_pulse_batch
    pb = make_pulse_batch(1024, 60.0, seed=1234)
    ei = gpu_knn(pb.x[:, :3].contiguous(), 8, pb.batch)
    N = pb.x.shape[0]
    src, dst = ei[0], ei[1]
    assert bool((dst[1:] >= dst[:-1]).all())
    assert bool((pb.batch[src] == pb.batch[dst]).all()) and bool((src != dst).all())
    deg = torch.bincount(dst, minlength=N)
    n_of = pb.n_pulses.long()[pb.batch]
    lo = torch.minimum(n_of, torch.tensor(9)) - 1
    assert bool((deg >= lo).all()) and bool((deg <= 9).all())
    d2 = ((pb.x[src, :3] - pb.x[dst, :3]) ** 2)
    d2 = (d2[:, 0] + d2[:, 1]) + d2[:, 2]
    same = dst[1:] == dst[:-1]
    assert bool((d2[1:][same] >= d2[:-1][same]).all())  # ascending distance per query
    # a sample of events against the oracle
    sel = pb.batch < 48
    want = oracle.knn_graph(pb.x[sel, :3], 8, pb.batch[sel])
    assert torch.equal(ei[:, dst < int(sel.sum())], want)


def test_unsorted_batch_is_rejected():
    import graphnet_b200 as g
    x = torch.randn(6, 3, device=dev())
    with pytest.raises(ValueError):
        g.knn_graph(x, 2, torch.tensor([0, 1, 0, 1, 1, 1], device=dev()))


def test_argument_errors_return_status():
    import graphnet_b200.ops as ops
    from graphnet_b200._lib import GraphnetB200Error
    x = torch.randn(6, 3, device=dev())
    ptr = torch.tensor([0, 6], dtype=torch.int32, device=dev())
    with pytest.raises(GraphnetB200Error, match="k="):
        ops.knn_graph_ell(x, 500, ptr, [0, 1, 2])
    with pytest.raises(GraphnetB200Error, match="column"):
        ops.knn_graph_ell(x, 2, ptr, [0, 1, 5])
