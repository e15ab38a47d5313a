"""The reference's own stored deployment (Backend/embedding/*.npy + Backend/metadata/*.json) as a relocatable fixture.

`reduce_from_reference()` (authoring container only) reads the five metadata files and `video_mapping.json`, drops
what this path never reads (bounding boxes, uuids, byte sizes), and rewrites the hard-coded `E:\\...\\Backend\\`
prefix to the placeholder `{ROOT}/` with forward slashes (the layout a Linux deployment of the reference has).
`materialise(layout, stores, root)` writes that deployment under `root` for the reference's services (authoring
container) and for the product's services (GPU box) alike.  Raw stores come from tests/golden/search_golden.npz, which
already carries the reference's .npy files byte for byte.
"""
import gzip
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LAYOUT_PATH = os.path.join(HERE, "golden", "ref_layout.json.gz")
SEARCH_GOLDEN = os.path.join(HERE, "golden", "search_golden.npz")
REF = "/root/reference/Backend"
WIN_PREFIX = "E:\\Đồ án tôt nghiệp\\source_code\\Backend\\"
VIDEOS = ["video_test_3", "video_test_4", "Violence", "Violence_2"]
# positional callbacks each strategy takes after its own arguments (Backend/query_strategies.py signatures)
STRATEGY_CALLBACKS = {
    "query_by_text_clip": ("search", "conf", "fmt"),
    "query_by_text_with_adaptive_threshold": ("search", "conf", "fmt"),
    "query_by_keyword": ("kw", "fmt"),
    "query_by_text_and_keyword": ("search", "conf", "kw", "fmt"),
    "query_by_object": ("fmt",),
    "query_by_text_and_object": ("search", "conf", "fmt"),
    "query_by_text_object_and_keyword": ("search", "conf", "kw", "fmt"),
}


def _relocatable(path):
    if isinstance(path, str) and path.startswith(WIN_PREFIX):
        return "{ROOT}/" + path[len(WIN_PREFIX):].replace("\\", "/")
    return path


def _reduce_record(rec):
    out = {"filepath": _relocatable(rec.get("filepath")), "tags": rec.get("tags", []),
           "video": _relocatable(rec.get("video")), "frameid": rec.get("frameid"), "frameidx": rec.get("frameidx")}
    meta = rec.get("metadata") or {}
    out["metadata"] = {k: meta[k] for k in ("caption", "width", "height") if k in meta}
    for field in ("text_detections", "object_detections"):
        dets = (rec.get(field) or {}).get("detections", [])
        out[field] = {"detections": [{"label": d.get("label", ""), "confidence": d.get("confidence", 0)} for d in dets]}
    if "embedding_model" in rec:
        out["embedding_model"] = rec["embedding_model"]
    return out


def reduce_from_reference():
    mapping = json.load(open(f"{REF}/metadata/video_mapping.json", encoding="utf-8"))
    layout = {"mapping": {v: {k: _relocatable(x) for k, x in e.items()} for v, e in mapping.items()}, "metadata": {}}
    for name in VIDEOS:
        data = json.load(open(f"{REF}/metadata/{name}_metadata.json", encoding="utf-8"))
        layout["metadata"][f"{name}_metadata.json"] = [_reduce_record(r) for r in data]
    data = json.load(open(f"{REF}/metadata/output_samples.json", encoding="utf-8"))
    layout["metadata"]["output_samples.json"] = [_reduce_record(r) for r in data]
    return layout


def save_layout(layout):
    with gzip.open(LAYOUT_PATH, "wt", encoding="utf-8") as f:
        json.dump(layout, f, ensure_ascii=False, separators=(",", ":"))


def load_layout():
    with gzip.open(LAYOUT_PATH, "rt", encoding="utf-8") as f:
        return json.load(f)


def _place(obj, root):
    if isinstance(obj, str):
        return obj.replace("{ROOT}", root)
    if isinstance(obj, list):
        return [_place(x, root) for x in obj]
    if isinstance(obj, dict):
        return {k: _place(v, root) for k, v in obj.items()}
    return obj


def unplace(obj, root):
    """Inverse of the placement: absolute paths under `root` back to `{ROOT}/...` (for comparing outputs)."""
    if isinstance(obj, str):
        return obj.replace(root, "{ROOT}")
    if isinstance(obj, list):
        return [unplace(x, root) for x in obj]
    if isinstance(obj, tuple):
        return [unplace(x, root) for x in obj]
    if isinstance(obj, dict):
        return {k: unplace(v, root) for k, v in obj.items()}
    return obj


def materialise(root, layout=None):
    """Writes the deployment under `root`; returns (base_dir, metadata_dir, embedding_dir)."""
    layout = layout or load_layout()
    root = str(root)
    base, meta, emb = (os.path.join(root, "static", "processed_frames"), os.path.join(root, "metadata"),
                       os.path.join(root, "embedding"))
    for d in (base, meta, emb):
        os.makedirs(d, exist_ok=True)
    with open(os.path.join(meta, "video_mapping.json"), "w", encoding="utf-8") as f:
        json.dump(_place(layout["mapping"], root), f, ensure_ascii=False, indent=4)
    for fname, records in layout["metadata"].items():
        with open(os.path.join(meta, fname), "w", encoding="utf-8") as f:
            json.dump(_place(records, root), f, ensure_ascii=False)
    blob = np.load(SEARCH_GOLDEN)
    for name in VIDEOS:
        np.save(os.path.join(emb, f"{name}_embeddings.npy"), blob[f"{name}/raw"])
    np.save(os.path.join(emb, "image_embeddings.npy"), blob["video_test_3/raw"])   # the legacy default store
    return base, meta, emb
