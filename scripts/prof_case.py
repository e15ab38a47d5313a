"""Small fixed workloads for ncu (one kernel family each; run plain first, then under ncu -- B200_PROFILING.md).
    python scripts/prof_case.py encode|encode1|text1|scan_big|scan_small|scan_q1 [steps]
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402
import bench  # noqa: E402

pkg = ge.load_package()

case = sys.argv[1]
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
if case == "encode":
    sd = pkg.clip.random_visual_state_dict("ViT-B/32", seed=1)
    enc = pkg.vit.VisionEncoder.from_state_dict(sd, "ViT-B/32", device=dev, max_batch=256, prefix="visual.")
    g = torch.Generator().manual_seed(0)
    frames = torch.randint(0, 256, (256, 224, 224, 3), generator=g, dtype=torch.uint8).to(dev)
    for _ in range(steps):
        out = enc.encode(frames, normalise=True)
    torch.cuda.synchronize()
    print("encode ok", float(out.float().abs().sum()))
elif case in ("encode1", "text1"):
    # single-frame / single-query latency case (EVR_NO_GRAPH=1 so that ncu sees every kernel)
    if case == "encode1":
        sd = pkg.clip.random_visual_state_dict("ViT-B/32", seed=1)
        enc = pkg.vit.VisionEncoder.from_state_dict(sd, "ViT-B/32", device=dev, max_batch=8, prefix="visual.")
        x = torch.randint(0, 256, (1, 224, 224, 3), dtype=torch.uint8).to(dev)
    else:
        enc = pkg.text.TextEncoder.from_state_dict(pkg.clip.random_text_state_dict("ViT-B/32", seed=2), device=dev,
                                                   max_batch=8)
        x = torch.zeros((1, 77), dtype=torch.int32, device=dev)
        x[:, 0], x[:, 1:9], x[:, 9] = 49406, 1234, 49407
    for _ in range(steps):
        out = enc.encode(x, normalise=True)
    torch.cuda.synchronize()
    print(case, "ok", float(out.float().abs().sum()))
else:
    if case.startswith("scan:"):      # scan:<rows>:<queries>
        _, n_s, q_s = case.split(":")
        custom = (int(n_s), int(q_s))
    else:
        custom = None
    n, q = custom or {"scan_c3": (10_000_000, 10_000), "scan_big": (2_000_000, 10_000), "scan_small": (10_000_000, 32), "scan_q1": (10_000_000, 1),
            "scan_q128": (10_000_000, 128)}[case]
    store = pkg.search.FrameStore(bench.make_store(torch, dev, n, seed=2))
    g = torch.Generator().manual_seed(3)
    qq = torch.randn((q, 512), generator=g)
    qq = (qq / qq.norm(dim=-1, keepdim=True)).to(dev).half()
    for _ in range(steps):
        sc, idx = store.search(qq, 100)
    torch.cuda.synchronize()
    print(case, "ok", float(sc.sum()))
