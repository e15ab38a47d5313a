"""Developer probe: small-batch latency with and without the captured-graph path (run twice, EVR_NO_GRAPH=0/1)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
dev = torch.device("cuda:0")
gen = torch.Generator(device=dev).manual_seed(0)
all_frames = torch.randint(0, 256, (256, 224, 224, 3), dtype=torch.uint8, device=dev, generator=gen)
for B in (32, 8, 1):
    sd = pkg.clip.random_visual_state_dict("ViT-B/32", seed=1)
    enc = pkg.vit.VisionEncoder.from_state_dict(sd, "ViT-B/32", device=dev, max_batch=B, prefix="visual.")
    frames = all_frames[:B]
    for _ in range(3):
        enc.encode(frames, normalise=True)
    torch.cuda.synchronize()

    def timed(fn, n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    n = 30 if B == 256 else 200
    ms = timed(lambda: enc.encode(frames, normalise=True), n)
    print(f"vit B={B}: {ms:.4f} ms/step (EVR_NO_GRAPH={os.environ.get('EVR_NO_GRAPH', '0')})", flush=True)
    del enc

tsd = pkg.clip.random_text_state_dict("ViT-B/32", seed=2)
tenc = pkg.text.TextEncoder.from_state_dict(tsd, device=dev, max_batch=32)
for B in (1, 8, 32):
    tok = torch.zeros((B, 77), dtype=torch.int32, device=dev)
    tok[:, 0], tok[:, 1:9], tok[:, 9] = 49406, 1234, 49407
    for _ in range(3):
        tenc.encode(tok, normalise=True)
    ms = timed(lambda: tenc.encode(tok, normalise=True), 200)
    print(f"text B={B}: {ms:.4f} ms/step", flush=True)
