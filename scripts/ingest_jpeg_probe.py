"""Developer probe / bench helper: folder ingest of 1280x720 JPEG key-frames end to end (files -> .npy), PIL decode on
host threads vs nvJPEG on the device.   python scripts/ingest_jpeg_probe.py [n_frames]"""
import json
import os
import shutil
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402


def run(n_frames=1024, quiet=False):
    pkg = ge.load_package()
    src = sorted(os.path.join(ROOT, "tests", "golden", "ref_frames", f)
                 for f in os.listdir(os.path.join(ROOT, "tests", "golden", "ref_frames")) if f.endswith(".jpg"))
    tmp = tempfile.mkdtemp(prefix="evr_ingest_")
    folder = os.path.join(tmp, "frames", "vid")
    os.makedirs(folder)
    os.makedirs(os.path.join(tmp, "metadata"))
    os.makedirs(os.path.join(tmp, "embedding"))
    nbytes = 0
    for i in range(n_frames):
        dst = os.path.join(folder, f"{i}.jpg")
        shutil.copyfile(src[i % len(src)], dst)
        nbytes += os.path.getsize(dst)
    sd = pkg.clip.random_visual_state_dict("ViT-B/32", seed=1)
    out = {"frames": n_frames, "jpeg_bytes": nbytes, "frame": "1280x720 (the reference's own key-frames, repeated)"}
    rows = {}
    for mode in (False, True):
        ps = pkg.services.PathService(os.path.join(tmp, "frames"), os.path.join(tmp, "metadata"),
                                      os.path.join(tmp, "embedding"))
        cs = pkg.services.CacheService()
        svc = pkg.services.EmbeddingService(cs, ps, pkg.services.DataService(ps, cs), "cuda",
                                            load_kwargs=dict(state_dict=sd, max_batch=256), gpu_jpeg=mode)
        svc.extract_and_save_embeddings_from_folder(folder, None, "warm")          # warm-up pass (pinned buffers, plans)
        t0 = time.perf_counter()
        path = svc.extract_and_save_embeddings_from_folder(folder, None, "vid")
        dt = time.perf_counter() - t0
        rows[mode] = np.load(path).astype(np.float32)
        key = "nvjpeg" if mode else "pil_threads"
        out[key] = {"frames_per_s": round(n_frames / dt, 1), "seconds": round(dt, 3)}
        if mode:
            out[key]["backend"] = svc._jpeg.backend
    a, b = rows[False], rows[True]
    cos = (a * b).sum(-1) / (np.linalg.norm(a, axis=-1) * np.linalg.norm(b, axis=-1))
    out["min_cosine_nvjpeg_vs_pil"] = round(float(cos.min()), 6)
    shutil.rmtree(tmp, ignore_errors=True)
    if not quiet:
        print(json.dumps(out))
    return out


if __name__ == "__main__":
    run(int(sys.argv[1]) if len(sys.argv) > 1 else 1024)
