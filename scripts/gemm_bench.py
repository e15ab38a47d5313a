"""Developer probe: throughput of the tcgen05 GEMM building block on the ViT shapes vs torch.matmul (cuBLAS).
    python scripts/gemm_bench.py [M ...]
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
lib, ptr, sp, check = pkg._lib.lib, pkg._lib.ptr, pkg._lib.current_stream_ptr, pkg._lib.check
dev = torch.device("cuda:0")
Ms = [int(x) for x in sys.argv[1:]] or [12800, 51200]
shapes = [("qkv", 2304, 768, 0), ("out", 768, 768, 2), ("fc", 3072, 768, 1), ("proj", 768, 3072, 2)]


def timeit(fn, n=30):
    for _ in range(5):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


for M in Ms:
    for name, N, K, epi in shapes:
        A = torch.randn(M, K, device=dev).bfloat16()
        W = (torch.randn(N, K, device=dev) * 0.05).bfloat16()
        bias = torch.randn(N, device=dev)
        out = torch.zeros(M, N, device=dev, dtype=torch.float32 if epi >= 2 else torch.bfloat16)
        ms = timeit(lambda: check(lib.evr_gemm_bf16(ptr(A), ptr(W), ptr(bias), ptr(out), M, N, K, epi, sp()), "gemm"))
        ms_t = timeit(lambda: torch.matmul(A, W.t()))
        fl = 2.0 * M * N * K
        print(f"M={M:6d} {name:5s} N={N:4d} K={K:4d} epi={epi}: ours {ms*1e3:7.1f} us {fl/ms/1e9:7.1f} TF/s | "
              f"cuBLAS (no epilogue) {ms_t*1e3:7.1f} us {fl/ms_t/1e9:7.1f} TF/s", flush=True)
