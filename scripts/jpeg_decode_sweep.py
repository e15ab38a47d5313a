"""Developer probe: nvJPEG batched decode alone (host bytes -> RGB in HBM), batch size x host threads.
   python scripts/jpeg_decode_sweep.py"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def one(batch, total=2048):
    import torch
    import __graft_entry__ as ge
    pkg = ge.load_package()
    d = os.path.join(ROOT, "tests", "golden", "ref_frames")
    blobs = [open(os.path.join(d, f), "rb").read() for f in sorted(os.listdir(d)) if f.endswith(".jpg")]
    blobs = (blobs * (total // len(blobs) + 1))[:total]
    dec = pkg.preprocess.JpegDecoder("cuda")
    size = dec.size(blobs[0])
    for _ in range(2):
        dec.decode(blobs[:batch], size)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(0, total - batch + 1, batch):
        dec.decode(blobs[i:i + batch], size)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    n = (total // batch) * batch
    print(json.dumps({"batch": batch, "threads": os.environ.get("EVR_JPEG_THREADS", "1"), "backend": dec.backend,
                      "frames_per_s": round(n / dt, 1), "mean_jpeg_kb": round(sum(map(len, blobs)) / len(blobs) / 1024, 1)}))


if __name__ == "__main__":
    if len(sys.argv) > 1:
        one(int(sys.argv[1]))
    else:
        for extra in ({}, {"EVR_JPEG_NO_POOL": "1"}):
            for batch in (32, 128, 256, 512):
                print(extra, flush=True)
                subprocess.run([sys.executable, __file__, str(batch)], env=dict(os.environ, **extra))
