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


def _random_ell(dev, g, B=37, k=8, stride=None):
    import graphnet_b200.ops as ops
    from graphnet_b200.synthetic import make_pulse_batch
    pb = make_pulse_batch(B, 20.0, seed=5)
    x = pb.x.to(dev)
    ptr, _ = ops.batch_to_ptr(pb.batch.to(dev), B)
    ell = ops.knn_graph_ell(x, k, ptr, [0, 1, 2], stride=stride)
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


def test_edge_grad_pq_gather_matches_scatter(dev):
    """dP / dQ by gathering through the graph grouped by source (gn_ell_reverse +
    gn_edge_grad_pq_gather) against torch's index_add_; HT-only gn_edge_build_h."""
    import graphnet_b200.ops as ops
    g = torch.Generator(device=dev).manual_seed(11)
    for B, k in ((12, 8), (3, 16), (40, 4)):
        N, ptr, ell = _random_ell(dev, g, B=B, k=k)
        KP = ell.ell_stride
        E = N * KP
        Ep = (E + 7) // 8 * 8
        Kp = 96
        P = (torch.randn((N, Kp), generator=g, device=dev) * 0.5).half()
        Q = (torch.randn((N, Kp), generator=g, device=dev) * 0.5).half()
        nbr = ell.col.reshape(-1).long()
        filled = nbr >= 0
        tgt = torch.arange(N, device=dev).repeat_interleave(KP)
        # reverse lists: every filled slot exactly once, grouped by source, ascending
        rstart, rcount, rslot = ops.ell_reverse(ell, ptr)
        assert int(rcount.sum()) == int(filled.sum())
        assert torch.equal(rcount.long(), torch.bincount(nbr[filled], minlength=N))
        rs, rc, sl = rstart.tolist(), rcount.tolist(), rslot.tolist()
        nb = nbr.tolist()
        for j in range(N):
            lst = sl[rs[j]:rs[j] + rc[j]]
            assert lst == sorted(lst) and all(nb[e] == j for e in lst)
        HT = torch.zeros((Kp, Ep), dtype=torch.bfloat16, device=dev)
        ops.edge_build_h(P, Q, Kp, ell, None, HT)
        pre = (P[tgt] + Q[nbr.clamp_min(0)])                     # fp16 sum, as the forward forms it
        Href = torch.relu(P.float()[tgt] + Q.float()[nbr.clamp_min(0)]) * filled[:, None]
        assert torch.equal(HT[:, :E], Href.to(torch.bfloat16).t())
        dH = torch.randn((E, Kp), generator=g, device=dev).to(torch.bfloat16)
        dPQ = torch.full((N, 2 * Kp), 9.0, dtype=torch.bfloat16, device=dev)
        dP32 = torch.empty((N, Kp), dtype=torch.float32, device=dev)
        dQ32 = torch.empty((N, Kp), dtype=torch.float32, device=dev)
        ops.edge_grad_pq_gather(dH, P, Q, Kp, ell, ptr, (rstart, rcount, rslot), dPQ[:, :Kp],
                                dPQ[:, Kp:], dP32, dQ32)
        m = dH.float() * (pre.float() > 0) * filled[:, None]
        dPref = m.reshape(N, KP, Kp).sum(1)
        dQref = torch.zeros((N, Kp), device=dev).index_add_(0, nbr.clamp_min(0), m)
        assert _rel(dP32, dPref) < 1e-6 and _rel(dQ32, dQref) < 1e-6
        assert torch.equal(dPQ[:, :Kp], dP32.to(torch.bfloat16))
        assert torch.equal(dPQ[:, Kp:], dQ32.to(torch.bfloat16))
        # without the fp32 copies, and twice: same bits (fixed summation order)
        again = torch.empty_like(dPQ)
        ops.edge_grad_pq_gather(dH, P, Q, Kp, ell, ptr, ops.ell_reverse(ell, ptr), again[:, :Kp],
                                again[:, Kp:])
        assert torch.equal(again, dPQ)
        # default form: one pass, dQ accumulated by vector reductions (order of additions not fixed)
        one = torch.full((N, 2 * Kp), 9.0, dtype=torch.bfloat16, device=dev)
        p32 = torch.empty((N, Kp), dtype=torch.float32, device=dev)
        q32 = torch.empty((N, Kp), dtype=torch.float32, device=dev)
        ops.edge_grad_pq_gather(dH, P, Q, Kp, ell, ptr, None, one[:, :Kp], one[:, Kp:], p32, q32)
        assert torch.equal(p32, dP32) and _rel(q32, dQref) < 1e-6
        assert torch.equal(one[:, :Kp], dPQ[:, :Kp])
        assert torch.equal(one[:, Kp:], q32.to(torch.bfloat16))


@pytest.mark.parametrize("k,a,b", [(8, 336, 256), (16, 128, 256), (4, 40, 48), (8, 62, 130)])
def test_edge_relu_bits_and_grad_z_mask(dev, k, a, b):
    """The training forward records [z > 0] of the second EdgeConv Linear as bits; the backward
    builds G2 from them.  Bits against z recomputed in fp64 from the same fp16 operands (they
    may differ only where z is within rounding of 0); G2 / G2T exactly dOut under those bits;
    the forward output is unchanged by recording."""
    import graphnet_b200.ops as ops
    from graphnet_b200.dynedge_bf16 import _slots, _round_up
    g = torch.Generator(device=dev).manual_seed(k + a)
    N, ptr, ell = _random_ell(dev, g, B=30, k=k, stride=_slots(k))
    KP = ell.ell_stride
    E = N * KP
    Ep = (E + 7) // 8 * 8
    Kp, b16 = _round_up(a + 2, 32), _round_up(b, 16)
    P = torch.zeros((N, Kp), device=dev)
    Q = torch.zeros((N, Kp), device=dev)
    P[:, :a] = torch.randn((N, a), generator=g, device=dev)
    P[:, a:a + 2] = 1.0
    Q[:, :a] = torch.randn((N, a), generator=g, device=dev)
    P, Q = P.half(), Q.half()
    W2p = torch.zeros((b16, Kp), device=dev)
    W2p[:b, :a + 2] = torch.randn((b, a + 2), generator=g, device=dev) / a ** 0.5
    W2p = W2p.half()
    out0 = torch.empty((N, b16), dtype=torch.bfloat16, device=dev)
    out1 = torch.empty((N, b16), dtype=torch.bfloat16, device=dev)
    ops.tc_edge_mlp_aggregate(P, Q, Kp, ell, KP, W2p, b, out0, None, [])
    mask = ops.edge_relu_mask(N, KP, dev)
    mask.fill_(-1)
    ops.tc_edge_mlp_aggregate(P, Q, Kp, ell, KP, W2p, b, out1, None, [], relu_mask=mask)
    assert torch.equal(out0[:, :b], out1[:, :b])
    # decode: slot e -> row (e // used) * 64 + e % used; channel c -> word c // 32, bit c % 32
    used = (64 // KP) * KP
    e = torch.arange(E, device=dev)
    rows = (e // used) * 64 + e % used
    c = torch.arange(b, device=dev)
    words = mask[(c // 32)[None, :], rows[:, None]]
    bits = ((words >> (c % 32)[None, :]) & 1).bool()
    nbr = ell.col.reshape(-1).long()
    filled = nbr >= 0
    tgt = torch.arange(N, device=dev).repeat_interleave(KP)
    h = torch.relu((P[tgt] + Q[nbr.clamp_min(0)]).double()) * filled[:, None]
    z = h @ W2p[:b].double().t()
    assert not bool(bits[~filled].any())
    wrong = bits != (z > 0)
    assert float(z[wrong].abs().max() if bool(wrong.any()) else 0.0) < 1e-3
    assert float(wrong.float().mean()) < 1e-3
    dOut = torch.randn((N, b16), generator=g, device=dev)[:, :b]   # a slice, as dskip's is
    G2 = torch.empty((E, b16), dtype=torch.bfloat16, device=dev)
    G2T = torch.zeros((b16, Ep), dtype=torch.bfloat16, device=dev)
    ops.edge_grad_z_mask(mask, dOut, b, ell, G2, G2T)
    G2ref = (dOut[tgt] * bits).to(torch.bfloat16)
    assert torch.equal(G2[:, :b], G2ref)
    assert torch.equal(G2T[:b, :E], G2ref.t())


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
