"""Calibration: cuBLAS (torch.matmul) on the GEMM shapes of the C2 forward, same strides, vs gn_tc_linear."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import graphnet_b200.ops as ops
dev = "cuda:0"
def timeit(fn, n=20):
    for _ in range(5): fn()
    s = [torch.cuda.Event(enable_timing=True) for _ in range(n)]; e = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
    for i in range(n):
        s[i].record(); fn(); e[i].record()
    torch.cuda.synchronize()
    t = sorted(a.elapsed_time(b) for a, b in zip(s, e))
    return t[len(t) // 2] * 1e3
for (M, N, K, dt, odt) in [(61443, 704, 256, torch.bfloat16, torch.float16), (61443, 336, 1088, torch.bfloat16, torch.bfloat16),
                           (61443, 256, 336, torch.bfloat16, torch.float32), (61443, 704, 72, torch.bfloat16, torch.float16)]:
    skip = torch.randn(M, 1088 + 64, device=dev).to(dt)
    A = skip[:, 64:64 + K]
    W = (torch.randn(N, K, device=dev) / K ** 0.5).to(dt)
    bias = torch.randn(N, device=dev)
    out = torch.empty(M, N, dtype=odt, device=dev)
    outb = torch.empty(M, N, dtype=dt, device=dev)
    Wt = W.t()
    t_cublas = timeit(lambda: torch.matmul(A, Wt, out=outb))
    t_addmm = timeit(lambda: torch.addmm(bias.to(dt), A, Wt, out=outb))
    t_ours = timeit(lambda: ops.tc_linear_bf16(A, W, bias, out, K=K, act="none"))
    print(f"M={M} N={N} K={K}: cuBLAS matmul {t_cublas:.1f} us, addmm(bias) {t_addmm:.1f} us, gn_tc_linear ({odt}) {t_ours:.1f} us")
