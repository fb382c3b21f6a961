"""Micro-benchmark of nc_gemm_bf16 on the hot-path GEMM shapes (CUDA events, L2 flushed)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nichecompass_b200 import ops

dev = "cuda"
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)


def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


shapes = [  # name, M, N, K, ta, tb, out dtype
    ("fc_l1 fwd", 8704, 1600, 5000, False, False, torch.bfloat16),
    ("lin_l fwd", 8704, 12800, 1600, False, False, torch.bfloat16),
    ("lin_l dX", 8704, 1600, 12800, False, True, torch.bfloat16),
    ("lin_l dW", 12800, 1600, 8704, True, True, torch.float32),
    ("fc_l1 dW", 1600, 5000, 8704, True, True, torch.float32),
    ("lin_r fwd", 1024, 12800, 1600, False, False, torch.bfloat16),
    ("lin_r dX", 1024, 1600, 12800, False, True, torch.float32),
    ("lin_r dW", 12800, 1600, 1024, True, True, torch.float32),
    ("fc_l1 fwd nb", 656, 1600, 5000, False, False, torch.bfloat16),
    ("lin_l fwd nb", 656, 12800, 1600, False, False, torch.bfloat16),
    ("lin_l dW nb", 12800, 1600, 656, True, True, torch.float32),
    ("decoder fwd", 64, 5000, 1600, False, False, torch.float32),
    ("decoder dW", 5000, 1600, 64, True, True, torch.float32),
    ("square 8k", 8192, 8192, 8192, False, False, torch.bfloat16),
]
for name, M, N, K, ta, tb, od in shapes:
    A = torch.randn((K, M) if ta else (M, K), device=dev).to(torch.bfloat16)
    B = torch.randn((K, N) if tb else (N, K), device=dev).to(torch.bfloat16)
    out = torch.empty(M, N, dtype=od, device=dev)
    ms = timeit(lambda: ops.gemm(A, B, ta, tb, out=out))
    At = A.t() if ta else A
    Bt = B if tb else B.t()
    ms_t = timeit(lambda: torch.matmul(At, Bt))
    fl = 2.0 * M * N * K
    print(f"{name:12s} M={M} N={N} K={K} ta={int(ta)} tb={int(tb)}: ours {ms:.3f} ms {fl/ms/1e9:.0f} TF/s | cuBLAS {ms_t:.3f} ms {fl/ms_t/1e9:.0f} TF/s", flush=True)
