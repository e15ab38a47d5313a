"""Golden for the ingest's host logic, produced by the reference ITSELF: the reference's unmodified
`EmbeddingService.extract_and_save_embeddings_from_folder` (services/embedding_service.py:425-536) runs over a folder of
its own key-frames plus the awkward entries (upper-case extension, a text file, a numeric and a non-numeric name) with
the weight-free `clip` stand-in of tests/refshim/ (so the rows depend only on pixels), on the relocated deployment of
tests/ref_layout.py.  Stored: the `.npy` it wrote (rows, dtype, file name relative to the deployment), the order of the
frames, the metadata file after the `embedding_model` stamp.  tests/test_ingest_reference.py runs the product's service
with the same stand-in on the CPU and must produce the same bytes.  Authoring container only:
    python tests/golden/make_ref_ingest_golden.py
"""
import contextlib
import io
import json
import os
import shutil
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
TESTS = os.path.dirname(HERE)
REF = "/root/reference/Backend"
for p in (REF, os.path.join(TESTS, "refshim"), TESTS):
    sys.path.insert(0, p)
import ref_layout  # noqa: E402

VIDEO = "Violence"


def make_folder(base):
    """frames of one video: the reference's key-frames + entries its listing must skip or keep"""
    folder = os.path.join(base, VIDEO)
    os.makedirs(folder, exist_ok=True)
    src = os.path.join(HERE, "ref_frames")
    names = sorted(f for f in os.listdir(src) if f.endswith(".jpg"))
    for n in names:
        shutil.copyfile(os.path.join(src, n), os.path.join(folder, n))
    shutil.copyfile(os.path.join(src, names[0]), os.path.join(folder, "12.jpg"))
    shutil.copyfile(os.path.join(src, names[1]), os.path.join(folder, "UPPER.JPG"))     # :452 is case-sensitive
    with open(os.path.join(folder, "notes.txt"), "w") as f:
        f.write("not a frame")
    return folder


def main():
    from services.cache_service import CacheService
    from services.data_service import DataService
    from services.embedding_service import EmbeddingService
    from services.path_service import PathService
    root = tempfile.mkdtemp(prefix="evr_ref_ingest_")
    base, meta, emb = ref_layout.materialise(root)
    folder = make_folder(base)
    ps, cs = PathService(base, meta, emb), CacheService()
    with contextlib.redirect_stdout(io.StringIO()):
        svc = EmbeddingService(cs, ps, DataService(ps, cs), "cpu")
        path = svc.extract_and_save_embeddings_from_folder(folder, None, VIDEO)
    rows = np.load(path)
    with open(ps.get_metadata_path(VIDEO), encoding="utf-8") as f:
        stamped = json.load(f)
    np.save(os.path.join(HERE, "ref_ingest_rows.npy"), rows)
    with open(os.path.join(HERE, "ref_ingest_golden.json"), "w", encoding="utf-8") as f:
        json.dump({"source": REF + "/services/embedding_service.py:425-536 (run unmodified, clip stand-in tests/refshim)",
                   "video": VIDEO, "npy": os.path.relpath(path, root), "dtype": str(rows.dtype), "shape": list(rows.shape),
                   "frames": sorted(f for f in os.listdir(folder) if f.endswith((".jpg", ".jpeg", ".png"))),
                   "active_model_after": svc.active_model,
                   "stamp": sorted({str(item.get("embedding_model")) for item in stamped}),
                   "metadata_items": len(stamped)}, f, indent=1)
    print(path, rows.shape, rows.dtype, float(np.linalg.norm(rows, axis=-1).min()))
    shutil.rmtree(root, ignore_errors=True)


if __name__ == "__main__":
    main()
