"""Developer probe: folder ingest (JPEG decode on host threads -> GPU resize/crop -> ViT-B/32 -> .npy) end to end,
serial decode (the reference's loop shape) vs threaded decode.   python scripts/ingest_probe.py [n_frames]"""
import importlib
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
services = importlib.import_module(pkg.__name__ + ".services")
from PIL import Image  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 600
tmp = tempfile.mkdtemp(prefix="ingest_", dir=os.path.join(ROOT, "gpurun_out"))
frames = os.path.join(tmp, "frames", "vid")
os.makedirs(frames)
rng = np.random.default_rng(0)
base = [Image.fromarray(rng.integers(0, 255, (45, 80, 3), dtype=np.uint8)).resize((1280, 720), Image.BICUBIC)
        for _ in range(8)]
for i in range(n):  # 1280x720 key-frames like the reference's fixtures (846 real 1280x720 frames)
    base[i % 8].save(os.path.join(frames, f"{i:06d}.jpg"), quality=90)
meta, emb = os.path.join(tmp, "metadata"), os.path.join(tmp, "embedding")
os.makedirs(meta), os.makedirs(emb)
json.dump({"vid": {"metadata_file": os.path.join(meta, "vid_metadata.json"),
                   "embeddings_file": os.path.join(emb, "vid_embeddings.npy"), "video_path": "",
                   "frames_dir": frames}}, open(os.path.join(meta, "video_mapping.json"), "w"))
json.dump([{"filepath": os.path.join(frames, f"{i:06d}.jpg"), "frameidx": i} for i in range(n)],
          open(os.path.join(meta, "vid_metadata.json"), "w"))
ps = services.PathService(os.path.join(tmp, "frames"), meta, emb)
cs = services.CacheService()
svc = services.EmbeddingService(cs, ps, services.DataService(ps, cs), "cuda",
                                load_kwargs=dict(random_init=True, max_batch=256))
out = {}
for workers in (1, None, 1, None):
    svc.decode_workers = workers
    t0 = time.perf_counter()
    path = svc.extract_and_save_embeddings_from_folder(frames, video_name="vid")
    dt = time.perf_counter() - t0
    e = np.load(path)
    out.setdefault(workers, e)
    print(f"decode_workers={workers}: {n} frames in {dt:.2f} s = {n / dt:.0f} frames/s  rows {e.shape}", flush=True)
print("identical rows:", np.array_equal(out[1], out[None]), "cores", os.cpu_count())
