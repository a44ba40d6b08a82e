#!/usr/bin/env python
"""Timeline of CTA 0 of the tcgen05 GEMM on the node-GEMM shape of C2 (profiling build).

    GN_B200_DEBUG_LIB=1 GN_TC_DEBUG=1 python tools/trace_linear.py [M N K]
"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import graphnet_b200.ops as ops  # noqa: E402
from graphnet_b200 import _lib  # noqa: E402

M, N, K = (int(v) for v in sys.argv[1:4]) if len(sys.argv) > 3 else (61443, 704, 256)
dev = "cuda:0"
# A is a column slice of a wider buffer, as in the forward (the skip-cat buffer: lda = 1088)
LDA = int(os.environ.get("GN_LDA", "1088"))
skip = torch.randn(M, max(LDA, K + 64), device=dev).to(torch.bfloat16)
A = skip[:, 64:64 + K] if skip.shape[1] >= K + 64 else skip[:, :K]
W = (torch.randn(N, K, device=dev) / K ** 0.5).to(torch.bfloat16)
bias = torch.randn(N, device=dev)
out = torch.empty(M, N, dtype=torch.float16, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
lib = _lib.load()
for _ in range(3):
    if not os.environ.get("GN_NO_FLUSH"):
        flush.fill_(1)
    else:
        A.mul_(1.0)   # the producer of A ran just before: A is dirty in L2
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    ops.tc_linear_bf16(A, W, bias, out, K=K, act="none")
    e.record()
    torch.cuda.synchronize()
print(f"M={M} N={N} K={K}: {s.elapsed_time(e) * 1e3:.1f} us")
buf = (ctypes.c_ulonglong * (10 * 16 * 4))()
lib.gn_debug_linear_trace(buf, 10 * 16 * 4)
tr = torch.tensor(list(buf), dtype=torch.int64).reshape(10, 16, 4)
t0 = int(tr[tr > 0].min())
names = ["TMA", "MMA"] + [f"epi{w}" for w in range(8)]
for item in range(16):
    if int(tr[0, item, 0]) == 0 and int(tr[1, item, 0]) == 0:
        break
    line = [f"item {item:2d}"]
    for w in (0, 1, 2, 6):
        v = [int(x) - t0 if int(x) > 0 else -1 for x in tr[w, item]]
        line.append(f"{names[w]} {v}")
    print("  ".join(line))

sb = (ctypes.c_ulonglong * (3 * 64))()
lib.gn_debug_linear_stages(sb, 3 * 64)
st = torch.tensor(list(sb), dtype=torch.int64).reshape(3, 64)
print("stage: TMA issue | full seen (latency) | commit | next issue into this slot - commit")
for g_ in range(40):
    if int(st[0, g_]) == 0:
        break
    iss, full, com = (int(st[k, g_]) - t0 for k in range(3))
    print(f"  g={g_:2d}  issue {iss:6d}  full {full:6d} (+{full - iss:5d})  commit {com:6d}")
