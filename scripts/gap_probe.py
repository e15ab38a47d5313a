"""Developer probe: busy time per kernel family and idle time between the kernels of one encode (device stamps).
   python scripts/gap_probe.py [ViT-B/32|ViT-L/14|ViT-B/16] [batch]"""
import collections
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
dev = torch.device("cuda:0")
name = sys.argv[1] if len(sys.argv) > 1 else "ViT-B/32"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 256
sd = pkg.clip.random_visual_state_dict(name, seed=1)
enc = pkg.vit.VisionEncoder.from_state_dict(sd, name, device=dev, max_batch=B, prefix="visual.")
frames = [torch.randint(0, 256, (B, 224, 224, 3), dtype=torch.uint8, device=dev) for _ in range(4)]
for i in range(40):
    enc.encode(frames[i % 4], normalise=True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
e0.record()
for i in range(40):
    enc.encode(frames[i % 4], normalise=True)
e1.record()
torch.cuda.synchronize()
print(f"{name} B={B}: {e0.elapsed_time(e1) / 40:.3f} ms per batch (40 batches, events)")
pkg._lib.profile_enable(2)
for i in range(8):
    enc.encode(frames[i % 4], normalise=True)
torch.cuda.synchronize()
tl = pkg._lib.profile_timeline()
pkg._lib.profile_enable(False)
gaps, dur = collections.defaultdict(list), collections.defaultdict(list)
for (f0, s0, t0), (f1, s1, t1) in zip(tl, tl[1:]):
    gaps[(f0, f1)].append((s1 - t0) * 1e-3)
for f, s, e in tl:
    dur[f].append((e - s) * 1e-3)
for k, v in sorted(gaps.items()):
    v = sorted(v)
    print(f"  {k[0]:10s} -> {k[1]:10s} n={len(v):4d}  median {v[len(v) // 2]:7.2f} us  mean {sum(v) / len(v):7.2f}")
for f, v in dur.items():
    print(f"  busy {f:10s} n={len(v):4d} mean {sum(v) / len(v):7.2f} us  total/batch {sum(v) / 8:8.1f} us")
