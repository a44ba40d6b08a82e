import torch, sys
sys.path.insert(0, '.')
import graphnet_b200.ops as ops
DEV='cuda:0'
for (M,N,K) in [(300,704,19),(1000,352,1088),(77,336,256),(4096,256,336),(129,320,24),(5,16,8)]:
    g = torch.Generator().manual_seed(M * 7 + N + K)
    Kp = (K + 7) // 8 * 8
    A = torch.randn(M, Kp + 16, generator=g).to(torch.bfloat16)
    W = (torch.randn(N, Kp, generator=g) / K ** 0.5).to(torch.bfloat16)
    bias = torch.randn(N, generator=g)
    out = torch.full((M, N + 8), -7.0, dtype=torch.float32, device=DEV)
    ops.tc_linear_bf16(A.to(DEV)[:, 8:], W.to(DEV), bias.to(DEV), out, K=K, act="none")
    want = A[:, 8:8 + K].float() @ W[:, :K].float().t() + bias
    got = out.cpu()[:, :N]
    err = (got - want).abs()
    print(M,N,K,"max err", float(err.max()), "at", divmod(int(err.argmax()), N))
    bad = err > 1e-3
    if bad.any():
        rows = bad.any(1).nonzero().flatten(); cols = bad.any(0).nonzero().flatten()
        print("  bad rows", rows[:5].tolist(), "..", rows[-3:].tolist(), len(rows), " bad cols", cols[:5].tolist(), "..", cols[-3:].tolist(), len(cols))
        # compare with using full Kp columns
        want2 = A[:, 8:8 + Kp].float() @ W[:, :Kp].float().t() + bias
        print("  err vs Kp-wide product", float((got - want2).abs().max()))
