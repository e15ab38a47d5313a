"""Developer probe (run under ncu --metrics gpu__time_duration.sum): the N=768 GEMMs with each epilogue."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
lib, ptr, sp, check = pkg._lib.lib, pkg._lib.ptr, pkg._lib.current_stream_ptr, pkg._lib.check
dev = torch.device("cuda:0")
M = 12800
import os
SHAPES = ((768, 768), (768, 3072), (2304, 768), (3072, 768))
for N, K in SHAPES:
    A = torch.randn(M, K, device=dev).bfloat16()
    W = (torch.randn(N, K, device=dev) * 0.05).bfloat16()
    bias = torch.randn(N, device=dev)
    for epi in ((0, 1, 3, 2) if not os.environ.get('EVR_PROBE_FAST') else (1,)):
        out = torch.zeros(M, N, device=dev, dtype=torch.float32 if epi >= 2 else torch.bfloat16)
        for _ in range(3):
            check(lib.evr_gemm_bf16(ptr(A), ptr(W), ptr(bias), ptr(out), M, N, K, epi, sp()), "gemm")
        torch.cuda.synchronize()
print("ok")
