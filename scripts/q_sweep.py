"""Developer probe: scan time vs query batch over a 10 M x 512 store (both roofline regimes and the ridge between)."""
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
for Q in (1, 32, 128, 129, 192, 256, 384, 512, 768, 1024, 2048, 4096, 10000):
    q = torch.randn((Q, 512), generator=g)
    q = (q / q.norm(dim=-1, keepdim=True)).to(dev).half()
    for _ in range(2):
        store.search(q, 100)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    n = 5 if Q <= 2048 else 2
    e0.record()
    for _ in range(n):
        store.search(q, 100)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    t_hbm = N * 512 * 2 / 6552e9 * 1e3
    t_tc = 2.0 * Q * N * 512 / 1371e12 * 1e3
    print(f"Q={Q:6d}: {ms:8.3f} ms  {Q / ms * 1e3:10.0f} q/s  roofline {max(t_hbm, t_tc):7.3f} ms  frac {max(t_hbm, t_tc) / ms:.3f}",
          flush=True)
