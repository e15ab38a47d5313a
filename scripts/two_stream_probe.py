"""Developer probe: does running two independent batches on two streams fill the GEMM wave tails?"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
clip = __import__("importlib").import_module(pkg.__name__ + ".clip")
dev = torch.device("cuda:0")
name, B = "ViT-B/32", 256
sd = {k[len("visual."):]: v for k, v in clip.random_visual_state_dict(name, 1).items() if k.startswith("visual.")}
encs = [pkg.vit.VisionEncoder.from_state_dict(sd, name, device=dev, max_batch=B) for _ in range(2)]
frames = [torch.randint(0, 256, (B, 224, 224, 3), dtype=torch.uint8, device=dev) for _ in range(8)]
streams = [torch.cuda.Stream(), torch.cuda.Stream()]


def run(n_streams, steps):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    if n_streams == 1:
        for i in range(steps):
            encs[0].encode(frames[i % 8], normalise=True)
    else:
        for st in streams:
            st.wait_stream(torch.cuda.current_stream())
        for i in range(steps):
            with torch.cuda.stream(streams[i & 1]):
                encs[i & 1].encode(frames[i % 8], normalise=True)
        for st in streams:
            torch.cuda.current_stream().wait_stream(st)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


for n in (1, 2, 1, 2):
    run(n, 6)
    ms = run(n, 40)
    print(f"{n} stream(s): {ms:.3f} ms per 256-frame batch, {B / ms * 1e3:.0f} frames/s", flush=True)
