"""Per-kernel-family time of one search call at several query-batch sizes (developer probe).
    python scripts/scan_breakdown.py [rows] [Q ...]
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402
import bench  # noqa: E402

pkg = ge.load_package()
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
qs = [int(x) for x in sys.argv[2:]] or [1, 8, 32, 128, 256, 1024]
dev = torch.device("cuda:0")
store = pkg.search.FrameStore(bench.make_store(torch, dev, rows, seed=2))
g = torch.Generator().manual_seed(3)
for Q in qs:
    q = torch.randn((Q, 512), generator=g)
    q = (q / q.norm(dim=-1, keepdim=True)).to(dev).half()
    for _ in range(3):
        store.search(q, 100)
    torch.cuda.synchronize()
    pkg._lib.profile_enable(True)
    n = 5
    for _ in range(n):
        store.search(q, 100)
    torch.cuda.synchronize()
    fam = pkg._lib.profile_read()
    pkg._lib.profile_enable(False)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        store.search(q, 100)
    e1.record()
    torch.cuda.synchronize()
    print(f"Q={Q:6d} call {e0.elapsed_time(e1) / n:8.3f} ms | scan {fam['topk_scan'][0] / n:8.3f} ms "
          f"({fam['topk_scan'][1] // n} launches) final {fam['topk_final'][0] / n:7.3f} ms", flush=True)
