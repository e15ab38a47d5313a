"""The ingest's host logic against the reference ITSELF.

tests/golden/ref_ingest_rows.npy / ref_ingest_golden.json hold what the reference's unmodified
`EmbeddingService.extract_and_save_embeddings_from_folder` wrote for a folder of its own key-frames (plus an upper-case
extension, a text file and a numeric name) when `clip` is the weight-free stand-in of tests/refshim/
(tests/golden/make_ref_ingest_golden.py).  The product's `EmbeddingService` takes the same stand-in as `clip_module` and
runs on the CPU here -- listing, order, batching, normalisation, file name, dtype, metadata stamp and the restored
active model are host logic -- and must write the same bytes.  Where the reference tree is present the two services
are also run side by side on a second folder (a frame that cannot be decoded, a video that is not in the mapping)."""
import contextlib
import importlib
import io
import json
import os
import shutil
import sys

import numpy as np
import pytest

import ref_layout
from conftest import PKG

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")
sys.path.insert(0, GOLDEN)
import make_ref_ingest_golden as gen  # noqa: E402  (folder recipe only; its main() needs the reference tree)


def _shim_clip():
    path = os.path.join(HERE, "refshim")
    if path not in sys.path:
        sys.path.insert(0, path)
    sys.modules.pop("clip", None)
    return importlib.import_module("clip")


def _product_service(root, base, meta, emb):
    S = importlib.import_module(PKG + ".services")
    ps, cs = S.PathService(base, meta, emb), S.CacheService()
    with contextlib.redirect_stdout(io.StringIO()):
        svc = S.EmbeddingService(cs, ps, S.DataService(ps, cs), "cpu", clip_module=_shim_clip(), gpu_preprocess=False)
    return svc, ps


def test_ingest_writes_what_the_reference_wrote(tmp_path):
    with open(os.path.join(GOLDEN, "ref_ingest_golden.json"), encoding="utf-8") as f:
        want = json.load(f)
    base, meta, emb = ref_layout.materialise(str(tmp_path))
    folder = gen.make_folder(base)
    svc, ps = _product_service(str(tmp_path), base, meta, emb)
    with contextlib.redirect_stdout(io.StringIO()):
        path = svc.extract_and_save_embeddings_from_folder(folder, None, want["video"])
    assert os.path.relpath(path, str(tmp_path)) == want["npy"]
    rows = np.load(path)
    ref = np.load(os.path.join(GOLDEN, "ref_ingest_rows.npy"))
    assert str(rows.dtype) == want["dtype"] and list(rows.shape) == want["shape"]
    assert np.array_equal(rows, ref)                       # bit for bit, row i = want["frames"][i]
    assert svc.active_model == want["active_model_after"]
    with open(ps.get_metadata_path(want["video"]), encoding="utf-8") as f:
        stamped = json.load(f)
    assert len(stamped) == want["metadata_items"]
    assert sorted({str(item.get("embedding_model")) for item in stamped}) == want["stamp"]


def test_ingest_side_by_side_with_the_reference(tmp_path):
    if not os.path.isdir(ref_layout.REF):
        pytest.skip("reference tree not present on this machine")
    for p in (ref_layout.REF, os.path.join(HERE, "refshim")):
        if p in sys.path:
            sys.path.remove(p)
        sys.path.insert(0, p)
    from services.cache_service import CacheService
    from services.data_service import DataService
    from services.embedding_service import EmbeddingService as RefEmbedding
    from services.path_service import PathService
    outs = {}
    for who in ("reference", "product"):
        root = tmp_path / who
        base, meta, emb = ref_layout.materialise(str(root))
        folder = gen.make_folder(base)
        with open(os.path.join(folder, "5.jpg"), "wb") as f:       # sorts after "12.jpg": cannot be decoded
            f.write(b"\xff\xd8 this is not a jpeg")
        if who == "reference":
            ps, cs = PathService(base, meta, emb), CacheService()
            with contextlib.redirect_stdout(io.StringIO()):
                svc = RefEmbedding(cs, ps, DataService(ps, cs), "cpu")
        else:
            svc, ps = _product_service(str(root), base, meta, emb)
        with contextlib.redirect_stdout(io.StringIO()):
            path = svc.extract_and_save_embeddings_from_folder(folder, "finetuned", "Violence")   # no checkpoint: falls back
        outs[who] = (np.load(path), os.path.relpath(path, str(root)), svc.active_model)
    (r_rows, r_rel, r_model), (p_rows, p_rel, p_model) = outs["reference"], outs["product"]
    assert r_rel == p_rel and r_model == p_model == "original"
    assert r_rows.shape == p_rows.shape == (10, 512)
    bad = 1                                                       # "12.jpg", "5.jpg", "Violence__..." in sorted order
    good = [i for i in range(10) if i != bad]
    assert np.array_equal(r_rows[good], p_rows[good])
    # the documented deviation: the reference's placeholder is zeros_like(previous image) -> 0/0 = NaN
    # after the normalisation; the product writes a zero row
    assert np.isnan(r_rows[bad]).all() and not p_rows[bad].any()
