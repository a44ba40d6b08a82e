"""GPU: per-event features, pooling, dense layers, EdgeConv, heads vs the oracle."""
import pytest
import torch

import oracle
from oracle import tasks as otasks
from conftest import load_golden, random_events

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _graph(ei, N):
    import graphnet_b200.ops as ops
    return ops.coo_to_csr(ei.to(DEV), N)


def _ptr(batch, B):
    import graphnet_b200.ops as ops
    ptr, flags = ops.batch_to_ptr(batch.to(DEV), B)
    assert int(flags.item()) == 0
    return ptr


@pytest.mark.parametrize("name", ["oscnext5", "upgrade5", "prometheus50", "adversarial"])
def test_global_variables_golden(name):
    import graphnet_b200.ops as ops
    g = load_golden(name)
    x = g["x"].to(DEV)
    N, F = x.shape
    B = g["n_pulses"].shape[0]
    gph = _graph(g["edge_index"], N)
    assert int(gph.flags.item()) == 0
    x0 = torch.full((N, 2 * F + 5 + 3), 7.0, device=DEV)
    gv = ops.global_variables(x, gph, _ptr(g["batch"], B), g["n_pulses"].to(DEV), x0=x0,
                              x0_cols=2 * F + 5 + 3)
    want = g["global_variables"]
    # homophily columns are ratios of exact integer counts: bit-exact
    assert torch.equal(gv.cpu()[:, F:F + 4], want[:, F:F + 4])
    assert torch.allclose(gv.cpu(), want, rtol=1e-5, atol=1e-6)
    x0 = x0.cpu()
    assert torch.equal(x0[:, :F], g["x"])
    assert torch.equal(x0[:, F:2 * F + 5], gv.cpu()[g["batch"]])
    assert bool((x0[:, 2 * F + 5:] == 0).all())


def test_global_variables_ell_graph_and_bf16_out():
    import graphnet_b200.ops as ops
    x, batch, n = random_events(9, 2, 50, seed=4)
    ptr = _ptr(batch, 9)
    gph = ops.knn_graph_ell(x.to(DEV), 8, ptr, [0, 1, 2])
    ei = oracle.knn_graph(x[:, :3], 8, batch)
    x0 = torch.empty((x.shape[0], 24), dtype=torch.bfloat16, device=DEV)
    gv = ops.global_variables(x.to(DEV), gph, ptr, n.to(DEV), x0=x0, x0_cols=24)
    want = oracle.global_variables(x, ei, batch, n)
    assert torch.allclose(gv.cpu(), want, rtol=1e-5, atol=1e-6)
    ref0 = torch.cat([x, want[batch]], 1).to(torch.bfloat16).float()
    assert torch.allclose(x0.float().cpu()[:, :19], ref0, rtol=1e-2, atol=1e-3)


@pytest.mark.parametrize("C", [256, 48, 7])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_segment_pool(C, dtype):
    import graphnet_b200.ops as ops
    g = torch.Generator().manual_seed(C)
    counts = torch.tensor([3, 0, 61, 1, 200, 17])
    batch = torch.repeat_interleave(torch.arange(6), counts)
    x = torch.randn(int(counts.sum()), C, generator=g).to(dtype)
    schemes = ["min", "max", "mean", "sum"]
    got = ops.segment_pool(x.to(DEV), _ptr(batch, 6), schemes).cpu()
    xf = x.float()
    want = torch.cat([oracle.scatter_min(xf, batch, 6)[0], oracle.scatter_max(xf, batch, 6)[0],
                      oracle.scatter_mean(xf, batch, 6), oracle.scatter_sum(xf, batch, 6)], 1)
    assert torch.equal(got[:, :2 * C], want[:, :2 * C])  # min/max are exact
    assert torch.allclose(got, want, rtol=1e-5, atol=1e-5)
    assert bool((got[1] == 0).all())  # empty event -> 0
    # strided input (column slice of a wider matrix) and scheme order
    wide = torch.randn(x.shape[0], C + 9).to(dtype)
    got2 = ops.segment_pool(wide.to(DEV)[:, 5:5 + C], _ptr(batch, 6), ["sum", "min"]).cpu()
    w2 = wide[:, 5:5 + C].float()
    assert torch.allclose(got2[:, :C], oracle.scatter_sum(w2, batch, 6), rtol=1e-5, atol=1e-5)
    assert torch.equal(got2[:, C:], oracle.scatter_min(w2, batch, 6)[0])


@pytest.mark.parametrize("M,N,K", [(1, 1, 1), (130, 336, 1043), (64, 64, 16), (1000, 128, 1024),
                                   (77, 3, 128), (513, 257, 19)])
@pytest.mark.parametrize("act", ["none", "relu", "gelu"])
def test_linear_f32(M, N, K, act):
    import graphnet_b200.ops as ops
    g = torch.Generator().manual_seed(M + N + K)
    A = torch.randn(M, K, generator=g)
    W = torch.randn(N, K, generator=g) / K ** 0.5
    b = torch.randn(N, generator=g)
    got = ops.linear_f32(A.to(DEV), W.to(DEV), b.to(DEV), act=act).cpu()
    want = torch.nn.functional.linear(A.double(), W.double(), b.double())
    want = {"none": lambda t: t, "relu": torch.relu, "gelu": torch.nn.functional.gelu}[act](want)
    assert torch.allclose(got.double(), want, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("M,N,K", [(300, 320, 19), (1001, 704, 19), (2049, 64, 7), (4096, 34, 32),
                                   (999, 320, 33), (515, 1024, 20), (700, 320, 36), (600, 128, 37)])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_linear_f32_16bit_output_small_k(M, N, K, dtype):
    """First EdgeConv node GEMM: fp32 inputs, K = 2F+5, [P | Q] written as 16-bit rows (a
    column slice of a wider buffer); the small-K kernel of csrc/dense_f32.cu and, where it does
    not apply (K > 32), the generic one."""
    import graphnet_b200.ops as ops
    g = torch.Generator().manual_seed(M + N + K)
    A = torch.randn(M, K + 3, generator=g)[:, :K]          # strided input rows
    W = torch.randn(N, K, generator=g) / K ** 0.5
    b = torch.randn(N, generator=g)
    wide = torch.full((M, N + 16), 7.0, dtype=dtype, device=DEV)
    out = wide[:, 8:8 + N]
    ops.linear_f32(A.to(DEV), W.to(DEV), b.to(DEV), act="none", out=out)
    want = torch.nn.functional.linear(A.double(), W.double(), b.double())
    tol = 2e-3 if dtype == torch.float16 else 1.6e-2
    assert torch.allclose(out.cpu().double(), want, rtol=tol, atol=tol)
    assert bool((wide[:, :8] == 7).all()) and bool((wide[:, 8 + N:] == 7).all())


def _mlp(W1, b1, W2, b2, act):
    f = {"relu": torch.relu, "gelu": torch.nn.functional.gelu}[act]

    def nn(t):
        h = f(torch.nn.functional.linear(t, W1, b1))
        if W2 is not None:
            h = f(torch.nn.functional.linear(h, W2, b2))
        return h
    return nn


@pytest.mark.parametrize("d,a,b", [(19, 128, 256), (256, 336, 256), (5, 17, 9), (33, 64, 130)])
@pytest.mark.parametrize("aggr", ["add", "mean", "max"])
def test_edgeconv_f32_csr(d, a, b, aggr):
    import graphnet_b200.ops as ops
    x, batch, _ = random_events(7, 1, 45, F=max(d, 3), seed=d + a)
    x = x[:, :d].contiguous()
    ei = oracle.knn_graph(x[:, :3], 8, batch)
    g = torch.Generator().manual_seed(a)
    W1 = torch.randn(a, 2 * d, generator=g) / (2 * d) ** 0.5
    b1 = torch.randn(a, generator=g) * 0.1
    W2 = torch.randn(b, a, generator=g) / a ** 0.5
    b2 = torch.randn(b, generator=g) * 0.1
    want = oracle.edge_conv(x.double(), ei, _mlp(W1.double(), b1.double(), W2.double(),
                                                 b2.double(), "relu"), aggr=aggr)
    Wpq, W2T = ops.edgeconv_prepare_f32(W1.to(DEV), W2.to(DEV))
    got = ops.edgeconv_f32(x.to(DEV), _graph(ei, x.shape[0]), Wpq, b1.to(DEV), W2T, b2.to(DEV),
                           act="relu", aggr=aggr).cpu()
    scale = want.abs().max()
    assert float((got.double() - want).abs().max() / scale) < 1e-5


def test_edgeconv_f32_ell_gelu_one_layer_and_views():
    import graphnet_b200.ops as ops
    x, batch, _ = random_events(5, 2, 30, F=12, seed=2)
    ptr = _ptr(batch, 5)
    gph = ops.knn_graph_ell(x.to(DEV), 4, ptr, [0, 1, 2])
    ei = oracle.knn_graph(x[:, :3], 4, batch)
    g = torch.Generator().manual_seed(1)
    W1 = torch.randn(40, 24, generator=g) / 5
    b1 = torch.randn(40, generator=g) * 0.1
    # one-layer MLP, gelu
    want = oracle.edge_conv(x.double(), ei, _mlp(W1.double(), b1.double(), None, None, "gelu"))
    Wpq, W2T = ops.edgeconv_prepare_f32(W1.to(DEV), None)
    assert W2T is None
    got = ops.edgeconv_f32(x.to(DEV), gph, Wpq, b1.to(DEV), None, None, act="gelu").cpu()
    assert torch.allclose(got.double(), want, rtol=1e-4, atol=1e-5)
    # two layers, input and output are column slices of wider buffers
    W2 = torch.randn(20, 40, generator=g) / 6
    b2 = torch.randn(20, generator=g) * 0.1
    want = oracle.edge_conv(x.double(), ei, _mlp(W1.double(), b1.double(), W2.double(),
                                                 b2.double(), "gelu"))
    Wpq, W2T = ops.edgeconv_prepare_f32(W1.to(DEV), W2.to(DEV))
    wide_in = torch.zeros(x.shape[0], 30, device=DEV)
    wide_in[:, 3:15] = x.to(DEV)
    wide_out = torch.full((x.shape[0], 50), -1.0, device=DEV)
    ops.edgeconv_f32(wide_in[:, 3:15], gph, Wpq, b1.to(DEV), W2T, b2.to(DEV), act="gelu",
                     out=wide_out[:, 10:30])
    assert torch.allclose(wide_out[:, 10:30].cpu().double(), want, rtol=1e-4, atol=1e-5)
    assert bool((wide_out[:, :10] == -1).all()) and bool((wide_out[:, 30:] == -1).all())


def test_coo_to_csr_flags_and_degree():
    import graphnet_b200.ops as ops
    ei = torch.tensor([[1, 2, 0, 3, 3], [0, 0, 2, 2, 2]])
    gph = ops.coo_to_csr(ei.to(DEV), 5)
    assert gph.rowptr.cpu().tolist() == [0, 2, 2, 5, 5, 5] and gph.col.cpu().tolist() == [1, 2, 0, 3, 3]
    assert int(gph.flags.item()) == 0 and int(gph.max_degree.item()) == 3
    bad = ops.coo_to_csr(torch.tensor([[1, 2], [2, 0]]).to(DEV), 3)
    assert int(bad.flags.item()) & 2
    oob = ops.coo_to_csr(torch.tensor([[1, 9], [0, 1]]).to(DEV), 3)
    assert int(oob.flags.item()) & 4
    empty = ops.coo_to_csr(torch.zeros(2, 0, dtype=torch.long).to(DEV), 4)
    assert empty.rowptr.cpu().tolist() == [0] * 5


@pytest.mark.parametrize("kind,p", [("EnergyReconstruction", 1), ("DirectionReconstructionWithKappa", 3),
                                    ("ZenithReconstruction", 1), ("ZenithReconstructionWithKappa", 2),
                                    ("AzimuthReconstructionWithKappa", 2)])
def test_task_heads(kind, p):
    import graphnet_b200.ops as ops
    g = torch.Generator().manual_seed(p)
    y = torch.randn(300, p, generator=g) * 30
    y[0] = 0.0
    y[1] = 1000.0  # softplus threshold branch
    got = ops.task_head(y.to(DEV), kind).cpu()
    want = otasks.head_transform(kind, y)
    assert got.shape == want.shape
    assert torch.allclose(got, want, rtol=2e-5, atol=1e-6)


def test_golden_heads_through_task_modules():
    import graphnet_b200 as gb
    for name in ("oscnext5", "prometheus50"):
        g = load_golden(name)
        cls = getattr(gb, g["head_kind"])
        t = cls(hidden_size=128).to(DEV)
        t._affine.weight.data.copy_(g["head_weight"])
        t._affine.bias.data.copy_(g["head_bias"])
        got = t(g["out"].to(DEV)).cpu()
        assert torch.allclose(got, g["head_out"], rtol=1e-4, atol=1e-5)


def test_eval_forward_runs_the_head_on_our_kernels():
    """An eval-mode StandardModel forward (no torch.no_grad() around it, as bench.py calls it) must
    run the task head through gn_linear_f32 + gn_task_head - no cuBLAS gemv, no eager softplus."""
    import graphnet_b200 as gb
    from graphnet_b200.synthetic import make_pulse_batch
    pb = make_pulse_batch(8, 20.0, seed=2)
    backbone = gb.DynEdge(7, global_pooling_schemes=["min", "max", "mean", "sum"], precision="bf16")
    task = gb.EnergyReconstruction(hidden_size=backbone.nb_outputs, target_labels="energy",
                                   loss_function=gb.LogCoshLoss())
    model = gb.StandardModel(tasks=[task], graph_definition=object(), gnn=backbone).to(DEV)
    model.inference()
    model.train(False)
    data = gb.Data(x=pb.x.to(DEV), batch=pb.batch.to(DEV), n_pulses=pb.n_pulses.to(DEV))
    model(data)   # warm-up (weight packing)
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        pred = model(data)[0]
        torch.cuda.synchronize()
    names = [e.key for e in prof.key_averages()]
    assert not pred.requires_grad
    if names:   # CUPTI available
        bad = [n for n in names if "gemv" in n.lower() or "softplus" in n.lower() or "gemm" in n.lower()]
        assert not bad, bad
        assert any("task_head_kernel" in n for n in names), names
