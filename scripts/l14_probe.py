"""Developer probe: ViT-L/14 (and B/16) encode throughput with the per-family kernel time split."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
clip = __import__("importlib").import_module(pkg.__name__ + ".clip")
dev = torch.device("cuda:0")
for name, B in (("ViT-L/14", 128), ("ViT-B/16", 256)):
    sd = {k[len("visual."):]: v for k, v in clip.random_visual_state_dict(name, 1).items() if k.startswith("visual.")}
    enc = pkg.vit.VisionEncoder.from_state_dict(sd, name, device=dev, max_batch=B)
    frames = torch.randint(0, 256, (B, 224, 224, 3), dtype=torch.uint8, device=dev)
    for _ in range(3):
        enc.encode(frames, normalise=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    n = 5
    for _ in range(n):
        enc.encode(frames, normalise=True)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    pkg._lib.profile_enable(True)
    enc.encode(frames, normalise=True)
    torch.cuda.synchronize()
    fam = {k: round(v[0], 3) for k, v in pkg._lib.profile_read().items() if v[1]}
    pkg._lib.profile_enable(False)
    gf = {"ViT-L/14": 162.03, "ViT-B/16": 35.1}[name]
    print(f"{name} B={B}: {ms:.2f} ms/step {B / ms * 1e3:.0f} frames/s {B * gf / ms:.0f} TF/s  families {fam}", flush=True)
    del enc
