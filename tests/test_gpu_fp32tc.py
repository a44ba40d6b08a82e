"""GPU: the fp32-accurate tensor-core kernels (fp16 / bf16 operand pairs, fp32 accumulation) against
float64 evaluations of the same inputs.  Tolerances: the split carries 22 bits (edge kernel) or 16
bits (GEMM operands), accumulation is fp32 - errors are those of an fp32 computation, 1e-5 and below
relative to the largest output."""
import pytest
import torch

import oracle
from conftest import random_events

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _w2_pair(W2p32):
    hi = W2p32.to(torch.float16)
    lo = (W2p32 - hi.float()).to(torch.float16)
    return hi.contiguous(), lo.contiguous()


@pytest.mark.parametrize("min_slots", [False, True])
@pytest.mark.parametrize("k,a,b", [(8, 336, 256), (8, 128, 256), (16, 336, 256), (4, 40, 48), (8, 62, 130)])
def test_tc_edge_mlp_aggregate_f32(k, a, b, min_slots):
    import graphnet_b200.ops as ops
    from graphnet_b200.dynedge_bf16 import _slots_min, _round_up
    from graphnet_b200.dynedge_bf16 import _slots as _slots_full
    _slots = _slots_min if min_slots else _slots_full
    x, batch, _ = random_events(40, 1, 70, seed=k + a)
    stack = torch.randn(40, x.shape[1])   # 40 coincident pulses: k+1 in-edges (overflow groups)
    stack[:, :3] = 0.125
    x = torch.cat([x, stack])
    batch = torch.cat([batch, torch.full((40,), 40)])
    N, B = x.shape[0], 41
    ptr, _ = ops.batch_to_ptr(batch.to(DEV), B)
    ell = ops.knn_graph_ell(x.to(DEV), k, ptr, [0, 1, 2])
    ei = oracle.knn_graph(x[:, :3], k, batch)
    Kp, b16 = _round_up(a + 2, 32), _round_up(b, 16)
    g = torch.Generator().manual_seed(a)
    P = torch.zeros(N, Kp)
    Q = torch.zeros(N, Kp)
    P[:, :a] = torch.randn(N, a, generator=g) * 3.0
    P[:, a] = 1.0
    Q[:, :a] = torch.randn(N, a, generator=g) * 3.0
    W2 = torch.zeros(b16, Kp)
    W2[:b, :a + 1] = torch.randn(b, a + 1, generator=g) / a ** 0.5
    hi, lo = _w2_pair(W2)
    out = torch.full((N, b + 4), -3.0, device=DEV)
    cols = [0, 1, 2] if b >= 200 else [1, 5, 33]
    coords = torch.empty(N, 3, device=DEV)
    ops.tc_edge_mlp_aggregate_f32(P.to(DEV), Q.to(DEV), Kp, ell, _slots(k), hi.to(DEV), lo.to(DEV), b,
                                  out, coords, cols)
    src, dst = ei[0], ei[1]
    h = torch.relu(P.double()[dst] + Q.double()[src])
    msg = torch.relu(h @ W2[:b].double().t())
    want = torch.zeros(N, b, dtype=torch.float64).index_add_(0, dst, msg)
    got = out.cpu()
    assert bool((got[:, b:] == -3.0).all())
    scale = float(want.abs().max())
    assert float((got[:, :b].double() - want).abs().max()) / scale < 2e-6
    assert float((coords.cpu().double() - want[:, cols]).abs().max()) / scale < 2e-6


def test_tc_edge_f32_csr_graph_with_large_degrees():
    """CSR graph with in-degrees far above the slot count: overflow groups, atomic fp32 adds."""
    import graphnet_b200.ops as ops
    g = torch.Generator().manual_seed(0)
    N, a, b, Kp = 300, 62, 64, 64
    dst = torch.sort(torch.randint(0, N, (6000,), generator=g)).values
    src = torch.randint(0, N, (6000,), generator=g)
    ei = torch.stack([src, dst])
    P = torch.zeros(N, Kp)
    Q = torch.zeros(N, Kp)
    P[:, :a] = torch.randn(N, a, generator=g)
    P[:, a] = 1.0
    Q[:, :a] = torch.randn(N, a, generator=g)
    W2 = torch.randn(b, Kp, generator=g) / 8
    hi, lo = _w2_pair(W2)
    out = torch.empty((N, b), device=DEV)
    ops.tc_edge_mlp_aggregate_f32(P.to(DEV), Q.to(DEV), Kp, ops.coo_to_csr(ei.to(DEV), N), 8,
                                  hi.to(DEV), lo.to(DEV), b, out, None, [])
    h = torch.relu(P.double()[dst] + Q.double()[src])
    want = torch.zeros(N, b, dtype=torch.float64).index_add_(0, dst, torch.relu(h @ W2.double().t()))
    assert float((out.cpu().double() - want).abs().max() / want.abs().max()) < 5e-6


@pytest.mark.parametrize("M,cols,kc", [(1000, 19, 24), (777, 256, 256), (33, 130, 136)])
def test_split_bf16_six_terms_is_fp32_grade(M, cols, kc):
    import graphnet_b200.ops as ops
    g = torch.Generator().manual_seed(M + 1)
    X = torch.randn(M, cols, generator=g) * 10.0
    A6 = ops.split_bf16(X.to(DEV), kc, terms=6)
    Nout = 72
    W = torch.randn(Nout, cols, generator=g) / cols ** 0.5
    w1 = W.to(torch.bfloat16)
    w2 = (W - w1.float()).to(torch.bfloat16)
    w3 = (W - w1.float() - w2.float()).to(torch.bfloat16)
    Wp = torch.zeros(Nout, 6 * kc, dtype=torch.bfloat16)
    for blk, piece in enumerate((w1, w2, w3, w1, w2, w1)):
        Wp[:, blk * kc:blk * kc + cols] = piece
    out = torch.empty(M, Nout, device=DEV)
    ops.tc_linear_bf16(A6, Wp.to(DEV), None, out, K=6 * kc, act="none")
    want = X.double() @ W.double().t()
    err = float((out.cpu().double() - want).abs().max() / want.abs().max())
    assert err < 4e-6, err   # (fp32 accumulation over K = 6 kc in the tensor core)


@pytest.mark.parametrize("M,cols,kc", [(1000, 19, 24), (777, 256, 256), (33, 130, 136)])
def test_split_bf16x3_and_the_gemm_it_feeds(M, cols, kc):
    import graphnet_b200.ops as ops
    g = torch.Generator().manual_seed(M)
    X = torch.randn(M, cols + 3, generator=g) * 10.0
    S = ops.split_bf16(X.to(DEV)[:, :cols], kc, terms=3)
    hi = X[:, :cols].to(torch.bfloat16)
    lo = (X[:, :cols] - hi.float()).to(torch.bfloat16)
    Sc = S.cpu()
    assert torch.equal(Sc[:, :cols], hi) and torch.equal(Sc[:, kc:kc + cols], lo)
    assert torch.equal(Sc[:, 2 * kc:2 * kc + cols], hi)
    assert bool((Sc[:, cols:kc] == 0).all()) and bool((Sc[:, kc + cols:2 * kc] == 0).all())
    # [hi | lo | hi] x [W_hi | W_hi | W_lo]^T on the bf16 tensor cores ~ fp32 product
    Nout = 72
    W = torch.randn(Nout, cols, generator=g) / cols ** 0.5
    Whi = W.to(torch.bfloat16)
    Wlo = (W - Whi.float()).to(torch.bfloat16)
    Wp = torch.zeros(Nout, 3 * kc, dtype=torch.bfloat16)
    Wp[:, :cols], Wp[:, kc:kc + cols], Wp[:, 2 * kc:2 * kc + cols] = Whi, Whi, Wlo
    out = torch.empty(M, Nout, device=DEV)
    ops.tc_linear_bf16(S, Wp.to(DEV), None, out, K=3 * kc, act="none")
    want = X[:, :cols].double() @ W.double().t()
    assert float((out.cpu().double() - want).abs().max() / want.abs().max()) < 3e-5


@pytest.mark.parametrize("with_edges", [True, False])
@pytest.mark.parametrize("kw", [
    dict(global_pooling_schemes=["min", "max", "mean", "sum"]),
    dict(global_pooling_schemes=["mean"], nb_neighbours=4, features_subset=[0, 1, 2, 3],
         dynedge_layer_sizes=[(64, 128), (168, 128)], post_processing_layer_sizes=[168, 128],
         readout_layer_sizes=[64]),
    dict(global_pooling_schemes=None, dynedge_layer_sizes=[(40, 50), (62, 100)],
         post_processing_layer_sizes=[96, 64, 33], readout_layer_sizes=[48]),
])
def test_fp32_forward_on_tensor_cores_matches_the_ffma_anchor(with_edges, kw):
    """DynEdge(precision="fp32") multiplies on the tensor cores by default; with
    fp32_tensor_cores=False the same model runs on the FFMA kernels.  Same graphs are required
    (kNN on latents that differ by 1e-6 could pick another neighbour at a near tie: events where
    that happens are compared at the looser bound)."""
    import graphnet_b200 as gb
    from graphnet_b200 import dynedge_x3
    x, batch, counts = random_events(37, 1, 90, seed=5, ties=False)
    torch.manual_seed(1)
    m = gb.DynEdge(7, precision="fp32", **kw).to(DEV).eval()
    assert dynedge_x3.eligible(m)
    ei = None
    if with_edges:
        ei = gb.knn_graph(x[:, :3].contiguous().to(DEV), kw.get("nb_neighbours", 8), batch.to(DEV),
                          num_events=37)
    data = gb.Data(x=x.to(DEV), edge_index=ei, batch=batch.to(DEV), n_pulses=counts.to(DEV))
    got = m(data)
    m.fp32_tensor_cores = False
    want = m(data)
    assert got.shape == want.shape
    if got.shape[0] == 37:     # per event
        err = (got - want).norm(dim=1) / want.norm(dim=1).clamp(min=1e-20)
        assert float(err.median()) < 3e-5, float(err.median())
        assert float((err < 1e-4).float().mean()) > 0.9, err
    else:                      # node-level output: rows of an event whose graph flipped differ
        err = (got - want).norm(dim=1) / want.norm(dim=1).clamp(min=1e-20)
        assert float(err.median()) < 3e-5, float(err.median())


def test_split_bf16_both_equals_the_two_separate_layouts():
    import graphnet_b200.ops as ops
    X = torch.randn(501, 130, generator=torch.Generator().manual_seed(4)).to(DEV) * 7.0
    kc = 136
    out3 = torch.zeros(501, 3 * kc + 8, dtype=torch.bfloat16, device=DEV)
    out6 = ops.split_bf16_both(X, kc, out3)
    assert torch.equal(out3[:, :3 * kc], ops.split_bf16(X, kc, terms=3))
    assert torch.equal(out6, ops.split_bf16(X, kc, terms=6))
    assert bool((out3[:, 3 * kc:] == 0).all())


def test_run_captured_replays_and_falls_back():
    """ops.run_captured: the fp32 forward as one updated CUDA graph gives the same bits as direct
    launches (GN_NO_GRAPH), for changing batch sizes; an exception inside the captured function
    leaves the library usable."""
    import os
    import graphnet_b200 as gb
    import graphnet_b200.ops as ops
    torch.manual_seed(3)
    m = gb.DynEdge(7, global_pooling_schemes=["min", "max", "mean", "sum"], precision="fp32").to(DEV).eval()
    datas = []
    for seed, (nev, lo, hi) in enumerate([(12, 5, 60), (1, 33, 33), (30, 1, 20)]):
        x, batch, counts = random_events(nev, lo, hi, seed=40 + seed)
        datas.append(gb.Data(x=x.to(DEV), batch=batch.to(DEV), n_pulses=counts.to(DEV)))
    got = [m(d) for d in datas + datas]
    os.environ["GN_NO_GRAPH"] = "1"
    try:
        want = [m(d) for d in datas + datas]
    finally:
        del os.environ["GN_NO_GRAPH"]
    for a, b in zip(got, want):
        assert torch.equal(a, b)

    def boom():
        ops.split_bf16(torch.ones(4, 8, device=DEV), 8)
        raise RuntimeError("inside the capture")
    with pytest.raises(RuntimeError, match="inside the capture"):
        ops.run_captured(12345, torch.device(DEV), boom)
    assert torch.equal(m(datas[0]), want[0])
