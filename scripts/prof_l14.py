"""ncu workload: two ViT-L/14 encodes of 64 frames (long-sequence attention_kernel, L = 257)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
dev = torch.device("cuda:0")
sd = pkg.clip.random_visual_state_dict("ViT-L/14", seed=1)
enc = pkg.vit.VisionEncoder.from_state_dict(sd, "ViT-L/14", device=dev, max_batch=64, prefix="visual.")
frames = torch.randint(0, 256, (64, 224, 224, 3), dtype=torch.uint8, device=dev)
for _ in range(2):
    out = enc.encode(frames, normalise=True)
torch.cuda.synchronize()
print("ok", float(out.abs().sum()))
