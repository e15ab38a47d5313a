"""Developer probe: evr_scene_deltas throughput on 1280x720 frames (HBM roofline: 3 bytes per pixel per frame)."""
import importlib
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
scenecut = importlib.import_module(pkg.__name__ + ".scenecut")

dev = torch.device("cuda:0")
T, H, W = 512, 720, 1280
frames = torch.randint(0, 256, (T, H, W, 3), dtype=torch.uint8, device=dev)
for down in (1, 2, 5):
    det = scenecut.ContentDetector(device=dev, downscale=down)
    for _ in range(3):
        det.delta_sums(frames)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    n = 10
    for _ in range(n):
        det.delta_sums(frames)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    gb = T * H * W * 3 / 1e9
    print(f"downscale {down}: {ms:.3f} ms for {T} frames = {T / ms * 1e3:.0f} frames/s, {gb / ms * 1e3:.0f} GB/s "
          f"of algorithmic frame bytes (each frame once)", flush=True)
host = frames[:9].cpu().numpy()
try:
    import cv2
    t0 = time.perf_counter()
    prev = None
    for f in host:
        hsv = cv2.cvtColor(f, cv2.COLOR_BGR2HSV).astype(np.int32)
        if prev is not None:
            [np.sum(np.abs(hsv[..., c] - prev[..., c])) for c in range(3)]
        prev = hsv
    dt = time.perf_counter() - t0
    print(f"cv2 + numpy (what ContentDetector runs): {8 / dt:.1f} frames/s")
except ImportError:
    pass
