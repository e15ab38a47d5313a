"""Developer probe: small-query-batch scan times with the strided sample pass vs the prefix sample (EVR_SAMPLE_PREFIX=1)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402
import bench  # noqa: E402

pkg = ge.load_package()
dev = torch.device("cuda:0")
N = 10_000_000
store = pkg.search.FrameStore(bench.make_store(torch, dev, N, seed=2))
g = torch.Generator().manual_seed(3)
qs = torch.randn((10000, 512), generator=g)
qs = (qs / qs.norm(dim=-1, keepdim=True)).to(dev).half()
for rnd in range(3):
    for mode in ("0", "1"):
        os.environ["EVR_SAMPLE_PREFIX"] = mode
        row = []
        for Q in (1, 8, 32, 128, 10000):
            q = qs[:Q].contiguous()
            for _ in range(3):
                store.search(q, 100)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            n = 10 if Q <= 128 else 3
            e0.record()
            for _ in range(n):
                store.search(q, 100)
            e1.record()
            torch.cuda.synchronize()
            row.append(f"Q={Q}: {e0.elapsed_time(e1) / n:.3f} ms")
        print("prefix" if mode == "1" else "strided", " | ".join(row), flush=True)
