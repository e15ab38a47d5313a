"""Runs the REFERENCE'S OWN service layer -- unmodified files imported from /root/reference/Backend -- over its own
stored deployment and records what it returns (first row).  Authoring container only:

    python tests/golden/make_ref_service_golden.py

Imported unmodified:  services/{path,cache,data,embedding,search}_service.py, query_strategies.py.
Executed unmodified:  `api_search` and the forwarding functions of app.py (:151-219, :381-613), compiled from the
                      file's own source text with the Flask decorator dropped (app.py itself cannot be imported: flask,
                      flask_cors, scenedetect, easyocr, ultralytics, umap are all absent).
Stand-ins on sys.path (tests/refshim): `clip` (weight-free deterministic encoders: the submodule is absent), `faiss`
(imported, never called), `unidecode`, `word_processing` (identity; its NLP / translation packages need the network).
Module constants that hold hard-coded E:\\ paths are re-pointed at the materialised layout (configuration, not code).

Outputs -> tests/golden/ref_layout.json.gz (the deployment) and tests/golden/ref_service_golden.json.gz:
  text_vectors                       query -> the [512] fp32 vector `encode_text` returned
  search_top_frames / confidence     every call the strategies made into EmbeddingService, with its answer
  keyword_frames                     SearchService.search_frames_by_keyword answers
  strategies                         the seven query_strategies functions' events
  semantic                           SearchService.search_semantic_with_clip events
  api                                api_search responses for a set of request bodies
"""
import ast
import gzip
import json
import os
import sys
import tempfile
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
TESTS = os.path.dirname(HERE)
sys.path.insert(0, TESTS)
import ref_layout  # noqa: E402

REF = ref_layout.REF
OUT = os.path.join(HERE, "ref_service_golden.json.gz")


def import_reference():
    sys.path.insert(0, REF)
    sys.path.insert(0, os.path.join(TESTS, "refshim"))   # ahead of Backend/: its word_processing.py needs pyvi etc.
    import clip  # the shim
    import query_strategies as ref_qs
    from services.cache_service import CacheService
    from services.data_service import DataService
    from services.embedding_service import EmbeddingService
    from services.path_service import PathService
    from services.search_service import SearchService
    return clip, ref_qs, dict(CacheService=CacheService, DataService=DataService, EmbeddingService=EmbeddingService,
                              PathService=PathService, SearchService=SearchService)


def reference_services(root, classes):
    base, meta, emb = ref_layout.materialise(root)
    ps = classes["PathService"](base, meta, emb)
    cs = classes["CacheService"]()
    ds = classes["DataService"](ps, cs)
    es = classes["EmbeddingService"](cs, ps, ds, "cpu")
    ss = classes["SearchService"](es, ds, ps, cs)
    return dict(path_service=ps, cache_service=cs, data_service=ds, embedding_service=es, search_service=ss)


def compile_app_functions(ref_qs, svc):
    """api_search + the forwarding functions, from app.py's own source (decorators dropped), over `svc`."""
    src = open(os.path.join(REF, "app.py"), encoding="utf-8").read()
    tree = ast.parse(src)
    wanted = {"api_search", "format_event_for_frontend", "search_frames_by_keyword", "extract_query_confidence",
              "search_top_frames", "search_top_frames_by_image", "search_by_image"}
    nodes = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in wanted]
    assert {n.name for n in nodes} == wanted
    for n in nodes:
        n.decorator_list = []
    module = ast.Module(body=nodes, type_ignores=[])
    import time
    from pathlib import Path
    request = types.SimpleNamespace(json=None)
    ns = dict(request=request, jsonify=lambda x: x, time=time, Path=Path,
              embedding_service=svc["embedding_service"], search_service=svc["search_service"],
              data_service=svc["data_service"], video_data_mapping=svc["path_service"].video_data_mapping)
    for name in ("query_by_text_clip", "query_by_text_with_adaptive_threshold", "query_by_keyword",
                 "query_by_text_and_keyword", "query_by_object", "query_by_text_and_object",
                 "query_by_text_object_and_keyword"):
        ns[name] = getattr(ref_qs, name)
    exec(compile(module, os.path.join(REF, "app.py"), "exec"), ns)
    return ns, request


def make_text_vectors(rng):
    """Deterministic 'text features' that produce a spread of cosines over the reference's stores: noisy copies of
    stored rows (high scores), mixtures (middling), a random vector (near zero)."""
    blob = np.load(ref_layout.SEARCH_GOLDEN)
    vectors = {}
    for name in ref_layout.VIDEOS:
        raw = blob[f"{name}/raw"].astype(np.float32)
        emb = raw / np.linalg.norm(raw, axis=-1, keepdims=True)
        n = emb.shape[0]
        a = emb[rng.integers(n)] + 0.02 * rng.standard_normal(512).astype(np.float32)
        rows = rng.choice(n, 3, replace=False)
        b = emb[rows].sum(0) + 0.03 * rng.standard_normal(512).astype(np.float32)
        vectors[f"a noisy copy of one frame of {name}"] = a * 7.5        # un-normalised, as encode_text emits
        vectors[f"trộn ba khung hình của {name}"] = b * 3.0               # diacritics in the cache key
    vectors["nothing in particular"] = rng.standard_normal(512).astype(np.float32)
    return {k: v.astype(np.float32) for k, v in vectors.items()}


def plain(obj):
    if isinstance(obj, dict):
        return {k: plain(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return [plain(v) for v in obj]
    if isinstance(obj, (np.floating,)):
        return float(obj)
    if isinstance(obj, (np.integer,)):
        return int(obj)
    return obj


def main():
    if not os.path.exists(ref_layout.LAYOUT_PATH) or "--relayout" in sys.argv:
        ref_layout.save_layout(ref_layout.reduce_from_reference())
        print("wrote", ref_layout.LAYOUT_PATH, os.path.getsize(ref_layout.LAYOUT_PATH))
    clip, ref_qs, classes = import_reference()
    rng = np.random.default_rng(20261105)
    vectors = make_text_vectors(rng)
    clip.TEXT_VECTORS.update(vectors)
    root = tempfile.mkdtemp(prefix="evr_ref_")
    svc = reference_services(root, classes)
    ref_qs.METADATA_DIR = os.path.join(root, "metadata")   # hard-coded E:\ constant (query_strategies.py:9) re-pointed
    es, ss, ds, ps = svc["embedding_service"], svc["search_service"], svc["data_service"], svc["path_service"]
    mapping = ps.video_data_mapping
    golden = {"text_vectors": {k: v.tolist() for k, v in vectors.items()}, "search_top_frames": [], "confidence": [],
              "keyword_frames": [], "strategies": [], "semantic": [], "api": [], "by_image": []}

    # ---- record every call the strategies make into the embedding service
    real_search, real_conf, real_kw = es.search_top_frames, es.extract_query_confidence, ss.search_frames_by_keyword

    def spy_search(query, top_k, video_name=None):
        out = real_search(query, top_k, video_name)
        golden["search_top_frames"].append({"query": query, "top_k": top_k, "video": video_name, "frames": list(out)})
        return out

    def spy_conf(frame_path, query, video_name=None):
        out = real_conf(frame_path, query, video_name)
        golden["confidence"].append({"frame": frame_path, "query": query, "video": video_name, "value": float(out)})
        return out

    def spy_kw(keyword, top_k, video_name=None):
        out = real_kw(keyword, top_k, video_name)
        golden["keyword_frames"].append({"keyword": keyword, "top_k": top_k, "video": video_name, "frames": list(out)})
        return out

    es.search_top_frames, es.extract_query_confidence, ss.search_frames_by_keyword = spy_search, spy_conf, spy_kw
    fmt = ds.format_event_for_frontend
    queries = list(vectors)

    # ---- direct service calls: scan results at several k (also k > N), default store, image-vector scan
    for video in [None] + ref_layout.VIDEOS:
        for q in queries[:3] + queries[-1:]:
            for k in (1, 5, 30, 1000):
                spy_search(q, k, video)
    blob = np.load(ref_layout.SEARCH_GOLDEN)
    for video in ref_layout.VIDEOS:
        feats = blob[f"{video}/queries"][:2].astype(np.float32)
        for j in range(feats.shape[0]):
            frames = es.search_top_frames_by_image(feats[j:j + 1], 12, video)
            golden["by_image"].append({"video": video, "features": feats[j].tolist(), "top_k": 12, "frames": list(frames)})
    for video in ref_layout.VIDEOS:
        frames = ds.load_frames_from_json(video)
        for f in (frames[0], frames[len(frames) // 2], "/somewhere/else/" + frames[-1], "missing.jpg"):
            spy_conf(f, queries[0], video)

    # ---- the seven strategies
    cbs = {"search": spy_search, "conf": spy_conf, "kw": spy_kw, "fmt": fmt}

    def run(name, args, **kwargs):
        callbacks = [cbs[c] for c in ref_layout.STRATEGY_CALLBACKS[name]]
        out = getattr(ref_qs, name)(*args, *callbacks, video_data_mapping=mapping, **kwargs)
        golden["strategies"].append({"fn": name, "args": list(args), "kwargs": kwargs,
                                     "events": ref_layout.unplace(plain(out), root)})

    for video in (None, "video_test_4", "Violence_2"):
        for q in ([f"a noisy copy of one frame of {video or 'video_test_3'}",
                   f"trộn ba khung hình của {video or 'video_test_3'}", "nothing in particular"]):
            for top_k, thr in ((5, 0.3), (10, 0.0)):
                run("query_by_text_clip", [q, top_k], video_name=video)
                run("query_by_text_with_adaptive_threshold", [q, thr, top_k], video_name=video)
                for kwd, tconf in (("Hotline", 0.5), ("đường", 0.2)):
                    run("query_by_text_and_keyword", [q, thr, top_k], keyword=kwd, text_confidence=tconf,
                        video_name=video)
                for obj, oconf in (("person", 0.4), ("MainData", 0.7), ("hotline", 0.3)):
                    run("query_by_text_and_object", [q, thr, top_k], object_keyword=obj, object_confidence=oconf,
                        video_name=video)
                    run("query_by_text_object_and_keyword", [q, thr, top_k], keyword="Hotline", text_confidence=0.5,
                        object_keyword=obj, object_confidence=oconf, video_name=video)
            run("query_by_text_and_keyword", [q, 0.0, 6], video_name=video)           # keyword / thresholds defaulted
            run("query_by_text_and_object", [q, 0.0, 6], video_name=video)
            run("query_by_text_object_and_keyword", [q, 0.0, 6], video_name=video)
        for kwd, thr in (("Hotline", 0.9), ("đường", 0.3), ("VIỆT NAM", 0.1), ("no such text", 0.0)):
            run("query_by_keyword", [kwd, thr, 7], video_name=video)
        for obj, thr in (("person", 0.5), ("car", 0.9), ("MainData", 0.7), ("hotline", 0.2), ("tivi", 0.1)):
            run("query_by_object", [obj, thr, 8], video_name=video)

    # ---- SearchService.search_semantic_with_clip (calls the embedding service directly, through the spies)
    for video in (None, "Violence"):
        for q in (f"a noisy copy of one frame of {video or 'video_test_3'}", "nothing in particular"):
            for thr in (0.0, 0.3):
                out = ss.search_semantic_with_clip(q, thr, 6, video)
                golden["semantic"].append({"query": q, "threshold": thr, "top_k": 6, "video": video,
                                           "events": ref_layout.unplace(plain(out), root)})

    # ---- the endpoint
    ns, request = compile_app_functions(ref_qs, svc)
    bodies = []
    q3, q4 = "a noisy copy of one frame of video_test_3", "a noisy copy of one frame of video_test_4"
    for method in ("text_clip", "text_adaptive", "keyword_only", "text_keyword", "object_only", "text_object",
                   "text_object_keyword", "something_else"):
        bodies.append({"query": q3, "search_method": method, "adaptive_threshold": 0.3, "keyword": "Hotline",
                       "object": "person", "top_k": 6})
        bodies.append({"query": q4, "search_method": method, "adaptive_threshold": 0.2, "text_confidence": 0.6,
                       "object_confidence": 0.45, "videoId": "video-2", "keyword": "đường", "object": "person"})
    bodies += [
        {"query": q3},                                                     # all defaults: method "text" -> adaptive 0.5
        {"query": ""},                                                     # empty query -> no dispatch
        {"query": q3, "search_method": "text_clip", "top_k": 3, "model": "finetuned"},   # model not loaded: stays original
        {"query": "trộn ba khung hình của Violence", "search_method": "text_clip", "videoId": "video-3", "top_k": 4},
        {"query": q3, "search_method": "text_object", "object": "person", "adaptive_threshold": 0.0,
         "enableClipSimilarity": True, "top_k": 5},
        {"query": q3, "search_method": "text_clip", "videoId": "video-9"},  # beyond the mapping: no video filter
        {"query": q3, "search_method": "text_clip", "videoId": "clip-1"},
        {"query": q3, "search_method": "text_clip", "videoId": "video-x"},  # int('x') -> 500
        {"search_type": "image", "query": q3},                             # image search without an image -> nothing
    ]
    for body in bodies:
        request.json = dict(body)
        resp = ns["api_search"]()
        payload, status = (resp if isinstance(resp, tuple) else (resp, 200))
        golden["api"].append({"body": body, "status": status, "response": ref_layout.unplace(plain(payload), root)})

    with gzip.open(OUT, "wt", encoding="utf-8") as f:
        json.dump(golden, f, ensure_ascii=False, separators=(",", ":"))
    print("wrote", OUT, os.path.getsize(OUT), {k: len(v) for k, v in golden.items()})


if __name__ == "__main__":
    main()
