"""Backward-pass building blocks (csrc/backward.cu, split-K tc_linear) against plain torch."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda:0")


def _rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))


@pytest.mark.parametrize("M,N,K,splits", [(256, 352, 5000, 0), (1000, 300, 256, 1), (704, 19, 3001, 0),
                                           (130, 64, 64, 8), (336, 1048, 61443, 0)])
def test_tc_linear_accumulate(dev, M, N, K, splits):
    import graphnet_b200.ops as ops
    g = torch.Generator(device=dev).manual_seed(1)
    Kp = (K + 7) // 8 * 8
    A = torch.randn((M, Kp), generator=g, device=dev).to(torch.bfloat16)
    W = torch.randn((N, Kp), generator=g, device=dev).to(torch.bfloat16)
    C0 = torch.randn((M, N), generator=g, device=dev)
    C = C0.clone()
    ops.tc_linear_accumulate(A, W, C, K, splits)
    ref = C0.double() + A[:, :K].double() @ W[:, :K].double().t()
    assert _rel(C, ref) < 1e-5


def test_transpose_colsum_relu_backward(dev):
    import graphnet_b200.ops as ops
    g = torch.Generator(device=dev).manual_seed(2)
    for rows, cols in [(1000, 336), (61, 19), (4097, 1048)]:
        X = torch.randn((rows, cols + 3), generator=g, device=dev).to(torch.bfloat16)[:, :cols]
        T = ops.transpose_16(X)
        assert torch.equal(T[:, :rows], X.t())
        s = torch.full((cols,), 2.0, device=dev)
        ops.colsum_accumulate(X, s)
        assert _rel(s, 2.0 + X.double().sum(0)) < 1e-5
        Xf = X.float() * 1.7
        s = torch.zeros(cols, device=dev)
        ops.colsum_accumulate(Xf, s)
        assert _rel(s, Xf.double().sum(0)) < 1e-5
        Y = torch.randn((rows, cols), generator=g, device=dev)
        Y[Y.abs() < 0.3] = 0
        for dy in (Xf, X):
            for y in (Y, Y.to(torch.bfloat16), None):
                out = ops.relu_backward(dy, y)
                ref = dy.float() if y is None else dy.float() * (y.float() > 0)
                assert torch.equal(out, ref.to(torch.bfloat16))


def _random_ell(dev, g, B=37, k=8):
    import graphnet_b200.ops as ops
    from graphnet_b200.synthetic import make_pulse_batch
    pb = make_pulse_batch(B, 20.0, seed=5)
    x = pb.x.to(dev)
    ptr, _ = ops.batch_to_ptr(pb.batch.to(dev), B)
    ell = ops.knn_graph_ell(x, k, ptr, [0, 1, 2])
    return x.shape[0], ptr, ell


def test_edge_backward_blocks(dev):
    import graphnet_b200.ops as ops
    g = torch.Generator(device=dev).manual_seed(3)
    N, ptr, ell = _random_ell(dev, g)
    KP = ell.ell_stride
    E = N * KP
    Ep = (E + 7) // 8 * 8
    Kp, b = 96, 64
    P = (torch.randn((N, Kp), generator=g, device=dev) * 0.5).half()
    Q = (torch.randn((N, Kp), generator=g, device=dev) * 0.5).half()
    nbr = ell.col.reshape(-1).long()
    filled = nbr >= 0
    tgt = torch.arange(N, device=dev).repeat_interleave(KP)
    Href = torch.relu(P.float()[tgt] + Q.float()[nbr.clamp_min(0)]) * filled[:, None]
    H = torch.empty((E, Kp), dtype=torch.bfloat16, device=dev)
    HT = torch.zeros((Kp, Ep), dtype=torch.bfloat16, device=dev)
    ops.edge_build_h(P, Q, Kp, ell, H, HT)
    assert torch.equal(H, Href.to(torch.bfloat16))
    assert torch.equal(HT[:, :E], H.t())

    Z = torch.randn((E, b), generator=g, device=dev).to(torch.bfloat16)
    dOut = torch.randn((N, b), generator=g, device=dev)
    G2 = torch.empty((E, b), dtype=torch.bfloat16, device=dev)
    G2T = torch.zeros((b, Ep), dtype=torch.bfloat16, device=dev)
    ops.edge_grad_z(Z, dOut, b, ell, G2, G2T)
    G2ref = (dOut[tgt] * (Z.float() > 0) * filled[:, None]).to(torch.bfloat16)
    assert torch.equal(G2, G2ref)
    assert torch.equal(G2T[:, :E], G2.t())

    dH = torch.randn((E, Kp), generator=g, device=dev).to(torch.bfloat16)
    dP = torch.empty((N, Kp), dtype=torch.bfloat16, device=dev)
    dQ = torch.zeros((N, Kp), dtype=torch.float32, device=dev)
    dP32 = torch.empty((N, Kp), dtype=torch.float32, device=dev)
    ops.edge_grad_pq(dH, H, Kp, ell, dP, dQ, dP32)
    m = dH.float() * (H.float() > 0) * filled[:, None]
    dPref = m.reshape(N, KP, Kp).sum(1)
    assert _rel(dP32, dPref) < 1e-5
    X = torch.randn((N, 19), generator=g, device=dev)
    acc = torch.ones((Kp, 19), device=dev)
    ops.outer_accumulate_f32(dP32, X, acc)
    assert _rel(acc, 1.0 + dP32.double().t() @ X.double()) < 1e-5
    dQref = torch.zeros((N, Kp), device=dev).index_add_(0, nbr.clamp_min(0), m)
    assert _rel(dP.float(), dPref) < 5e-3
    assert _rel(dQ, dQref) < 1e-5


def test_segment_pool_backward(dev):
    import graphnet_b200.ops as ops
    g = torch.Generator(device=dev).manual_seed(4)
    N, ptr, _ = _random_ell(dev, g)
    B = ptr.numel() - 1
    C = 200
    x = torch.relu(torch.randn((N, C), generator=g, device=dev))  # many exact zeros: ties for min
    schemes = ["min", "max", "mean", "sum"]
    pooled = ops.segment_pool(x, ptr, schemes)
    dp = torch.randn((B, C * 4), generator=g, device=dev)
    dX = ops.segment_pool_backward(x, ptr, schemes, pooled, dp)
    # reference: first row attaining min / max gets the gradient
    ref = torch.zeros_like(x)
    p = ptr.tolist()
    for e in range(B):
        xs = x[p[e]:p[e + 1]]
        n = xs.shape[0]
        if n == 0:
            continue
        r = ref[p[e]:p[e + 1]]
        imin = (xs == xs.min(0).values).float().argmax(0)
        imax = (xs == xs.max(0).values).float().argmax(0)
        cols = torch.arange(C, device=dev)
        r[imin, cols] += dp[e, 0:C]
        r[imax, cols] += dp[e, C:2 * C]
        r += dp[e, 2 * C:3 * C] / n
        r += dp[e, 3 * C:4 * C]
    assert _rel(dX, ref) < 1e-6
