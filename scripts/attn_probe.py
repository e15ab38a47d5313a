"""Developer probe: attention core, warp-level MMA kernels (impl 1) vs tcgen05 / TMEM (impl 2), same inputs."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
lib, check, ptr, stream = pkg._lib.lib, pkg._lib.check, pkg._lib.ptr, pkg._lib.current_stream_ptr
dev = torch.device("cuda:0")
for name, B, L, d, heads, causal in (("ViT-B/32", 256, 50, 768, 12, 0), ("ViT-L/14", 128, 257, 1024, 16, 0), ("ViT-B/16", 256, 197, 768, 12, 0),
                                     ("text", 64, 77, 512, 8, 1), ("text-1", 1, 77, 512, 8, 1)):
    bufs = [torch.randn((B * L, 3 * d), device=dev).to(torch.bfloat16) for _ in range(3)]   # rotate: inputs not in L2
    out = torch.empty((B * L, d), dtype=torch.bfloat16, device=dev)
    row = []
    for impl in (1, 2):
        for i in range(3):
            check(lib.evr_attention(ptr(bufs[i]), ptr(out), B, L, d, heads, causal, impl, stream()), "attn")
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 30
        torch.cuda.synchronize()
        e0.record()
        for i in range(n):
            check(lib.evr_attention(ptr(bufs[i % 3]), ptr(out), B, L, d, heads, causal, impl, stream()), "attn")
        e1.record()
        torch.cuda.synchronize()
        row.append(e0.elapsed_time(e1) * 1e3 / n)
    nbytes = B * L * d * 2 * 4
    print(f"{name:9s} B={B:3d} L={L}: mma.sync {row[0]:7.1f} us  tcgen05 {row[1]:7.1f} us  ({row[0] / row[1]:.2f}x)  "
          f"tcgen05 moves {nbytes / row[1] / 1e3:.0f} GB/s", flush=True)
