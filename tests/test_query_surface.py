"""The query surface above the scan: seven strategies, SearchService, api_search.

Pinned to the reference ITSELF: tests/golden/ref_service_golden.json.gz holds what the reference's unmodified
query_strategies.py / services/*.py / app.py:api_search returned on its own stored deployment
(tests/golden/make_ref_service_golden.py).  Here the product's host logic runs over the same deployment with the two
CLIP callbacks answered from the recording (so no GPU is needed) and must return the same events, value for value.
Where /root/reference is present (authoring container) the product is also fuzzed live against the reference's
functions on synthetic metadata with the awkward cases (missing fields, empty lists, ties, accents).
The GPU twin -- the same requests through the real EmbeddingService on the B200 -- is tests/test_query_surface_gpu.py.
"""
import gzip
import importlib
import json
import os
import random
import sys

import pytest

import ref_layout
from conftest import PKG

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_service_golden.json.gz")


@pytest.fixture(scope="module")
def golden():
    with gzip.open(GOLDEN, "rt", encoding="utf-8") as f:
        return json.load(f)


@pytest.fixture(scope="module")
def mods():
    return {m: importlib.import_module(f"{PKG}.{m}") for m in ("services", "query_strategies", "app")}


class ReplayEmbeddingService:
    """Answers the two CLIP callbacks from the reference's recorded answers (same arguments -> same answer)."""

    def __init__(self, golden):
        self.searches = {(r["query"], r["top_k"], r["video"]): r["frames"] for r in golden["search_top_frames"]}
        self.confidences = {(r["frame"], r["query"], r["video"]): r["value"] for r in golden["confidence"]}
        self.active_model, self.finetuned_model, self.preprocess, self.device = "original", None, None, "cpu"

    def get_active_model_name(self):
        return self.active_model

    def set_active_model(self, name):
        if name == "original":
            self.active_model = name
            return True
        return False   # the recording was made without a fine-tuned checkpoint

    def search_top_frames(self, query, top_k, video_name=None):
        return list(self.searches[(query, top_k, video_name)])

    def extract_query_confidence(self, frame_path, query, video_name=None):
        return self.confidences[(frame_path, query, video_name)]


@pytest.fixture(scope="module")
def deployment(tmp_path_factory, golden, mods):
    root = str(tmp_path_factory.mktemp("deploy"))
    base, meta, emb = ref_layout.materialise(root)
    S = mods["services"]
    ps, cs = S.PathService(base, meta, emb), S.CacheService()
    ds = S.DataService(ps, cs)
    es = ReplayEmbeddingService(golden)
    ss = S.SearchService(es, ds, ps, cs)
    mods["query_strategies"].METADATA_DIR = meta
    return dict(root=root, path_service=ps, cache_service=cs, data_service=ds, embedding_service=es,
                search_service=ss)


def test_keyword_frame_lookup_matches_reference(golden, deployment):
    ss = deployment["search_service"]
    assert len(golden["keyword_frames"]) > 50
    for r in golden["keyword_frames"]:
        assert ss.search_frames_by_keyword(r["keyword"], r["top_k"], r["video"]) == r["frames"], r["keyword"]


def test_seven_strategies_match_reference(golden, deployment, mods):
    qs, d = mods["query_strategies"], deployment
    cbs = {"search": d["embedding_service"].search_top_frames, "conf": d["embedding_service"].extract_query_confidence,
           "kw": d["search_service"].search_frames_by_keyword, "fmt": d["data_service"].format_event_for_frontend}
    seen = set()
    for r in golden["strategies"]:
        callbacks = [cbs[c] for c in ref_layout.STRATEGY_CALLBACKS[r["fn"]]]
        out = getattr(qs, r["fn"])(*r["args"], *callbacks, video_data_mapping=d["path_service"].video_data_mapping,
                                   **r["kwargs"])
        assert ref_layout.unplace(out, d["root"]) == r["events"], (r["fn"], r["args"], r["kwargs"])
        seen.add((r["fn"], bool(r["events"])))
    assert {fn for fn, nonempty in seen if nonempty} == set(ref_layout.STRATEGY_CALLBACKS)  # every strategy exercised


def test_semantic_search_matches_reference(golden, deployment):
    ss = deployment["search_service"]
    for r in golden["semantic"]:
        out = ss.search_semantic_with_clip(r["query"], r["threshold"], r["top_k"], r["video"])
        assert ref_layout.unplace(out, deployment["root"]) == r["events"]


def test_api_search_matches_reference(golden, deployment, mods):
    services = {k: deployment[k] for k in ("path_service", "cache_service", "data_service", "embedding_service",
                                           "search_service")}
    statuses = set()
    for r in golden["api"]:
        payload, status = mods["app"].api_search(dict(r["body"]), services)
        assert status == r["status"], r["body"]
        assert ref_layout.unplace(payload, deployment["root"]) == r["response"], r["body"]
        statuses.add(status)
    assert statuses == {200, 500}
    methods = {r["body"].get("search_method") for r in golden["api"] if r["response"].get("events")}
    assert {"text_clip", "text_adaptive", "keyword_only", "object_only", "text_object"} <= methods


def test_search_service_signatures_are_the_references(mods):
    import inspect
    S = mods["services"].SearchService
    assert list(inspect.signature(S.__init__).parameters)[1:] == ["embedding_service", "data_service", "path_service",
                                                                  "cache_service"]
    params = inspect.signature(S.search_by_image).parameters
    assert list(params)[:6] == ["self", "image_url", "adaptive_threshold", "top_k", "video_name", "preprocess"]
    assert params["video_name"].default is None and params["preprocess"].default is None
    assert list(inspect.signature(S.search_semantic_with_clip).parameters)[1:] == ["query", "adaptive_threshold",
                                                                                   "top_k", "video_name"]
    qs = mods["query_strategies"]
    assert list(inspect.signature(qs.query_by_text_object_and_keyword).parameters) == [
        "query", "adaptive_threshold", "top_k", "search_top_frames", "extract_query_confidence",
        "search_frames_by_keyword", "format_event_for_frontend", "keyword", "text_confidence", "object_keyword",
        "object_confidence", "video_name", "video_data_mapping"]


def test_unaccent_rule(mods):
    un = importlib.import_module(PKG + "._unaccent").unaccent
    assert un("Đường bộ Việt Nam") == "Duong bo Viet Nam"
    assert un("LO NGẠI HOẠT ĐỘNG CỦA THIẾT BỊ BAY KHÔNG NGƯỜI LÁI") == "LO NGAI HOAT DONG CUA THIET BI BAY KHONG NGUOI LAI"
    assert un("plain ascii 0889 24 24 24") == "plain ascii 0889 24 24 24"


# ---- live fuzz against the reference's own functions (authoring container only) ---------------------------------
def _reference_modules():
    if not os.path.isdir(ref_layout.REF):
        pytest.skip("reference tree not present on this machine")
    for p in (ref_layout.REF, os.path.join(os.path.dirname(os.path.abspath(__file__)), "refshim")):
        if p in sys.path:
            sys.path.remove(p)
        sys.path.insert(0, p)
    import query_strategies as ref_qs
    from services.data_service import DataService as RefData
    from services.search_service import SearchService as RefSearch
    return ref_qs, RefData, RefSearch


def _fuzz_metadata(rng, n):
    labels = ["Hotline", "Đường bộ Việt Nam", "duong", "person", "xe máy", "", "VIỆT NAM hotline", "tivi 24", "Person A"]
    objs = ["person", "car", "tv", "motorcycle", "xe", ""]
    records = []
    for i in range(n):
        rec = {"filepath": f"/frames/v/{1000 + i}.jpg", "frameid": f"{1000 + i}.jpg", "frameidx": 1000 + i,
               "video": rng.choice(["/videos/v.mp4", ""]), "tags": rng.choice([["MainData"], [], ["person", 7]])}
        if rng.random() < 0.8:
            rec["text_detections"] = {"detections": [
                {"label": rng.choice(labels), "confidence": rng.choice([0, 0.3, 0.5, 0.5, 0.92, 1.0])}
                for _ in range(rng.randrange(0, 4))]}
        if rng.random() < 0.7:
            rec["object_detections"] = {"detections": [
                {"label": rng.choice(objs), "confidence": rng.choice([0.0, 0.41, 0.65, 0.65, 0.88])}
                for _ in range(rng.randrange(0, 3))]}
        if rng.random() < 0.3:
            rec["metadata"] = {"caption": rng.choice(["a person on a xe máy", "nothing", ""])}
        if rng.random() < 0.1:
            rec["text_detections"] = {}
        if rng.random() < 0.05:
            rec["frameid"] = f"{1000 + i - 1}.jpg"   # duplicate frame id: first record wins
        records.append(rec)
    return records


def test_live_fuzz_against_reference(tmp_path, mods):
    ref_qs, RefData, RefSearch = _reference_modules()
    S, qs = mods["services"], mods["query_strategies"]
    rng = random.Random(7)
    meta = tmp_path / "metadata"
    meta.mkdir()
    for trial in range(6):
        records = _fuzz_metadata(rng, 60)
        (meta / "v_metadata.json").write_text(json.dumps(records, ensure_ascii=False), encoding="utf-8")
        mapping = {"v": {"metadata_file": str(meta / "v_metadata.json"), "embeddings_file": "", "video_path": "",
                         "frames_dir": ""}}
        (meta / "video_mapping.json").write_text(json.dumps(mapping))
        os.utime(meta / "v_metadata.json", (trial + 1, trial + 1))   # distinct mtime per trial (mtime-keyed index)
        ps, cs = S.PathService(str(tmp_path), str(meta), str(tmp_path)), S.CacheService()
        ours_ds, ref_ds = S.DataService(ps, cs), RefData(ps, cs)
        ours_ss, ref_ss = S.SearchService(None, ours_ds, ps, cs), RefSearch(None, ref_ds, ps, cs)
        names = [r["frameid"] for r in records]
        scores = {n: rng.choice([0.1, 0.25, 0.25, 0.31, 0.5, 0.77]) for n in names}

        def search(text, k, video=None):
            ranked = sorted(names, key=lambda n: -scores[n])[:k]
            return ranked + ["not-a-number.jpg"][:trial % 2]      # odd trials: a frame name int() cannot parse

        def conf(frame, text, video=None):
            return scores.get(frame, 0.0)

        for r in records[:25]:
            r2 = dict(r, clip_similarity=rng.choice([None, 0.0, 0.4, 0.95]))
            assert ours_ds.format_event_for_frontend(r2) == ref_ds.format_event_for_frontend(r2)
        for kw in ("hotline", "Đường", "duong", "24", "zzz"):
            assert ours_ss.search_frames_by_keyword(kw, 9, "v") == ref_ss.search_frames_by_keyword(kw, 9, "v")
        for fn, cb_names in ref_layout.STRATEGY_CALLBACKS.items():
            for thr in (0.0, 0.3, 0.5, 0.7):
                for extra in ({}, {"keyword": "Hotline"}, {"object_keyword": "person", "object_confidence": 0.6},
                              {"keyword": "đường", "text_confidence": 0.4, "object_keyword": "xe"}):
                    import inspect
                    allowed = set(inspect.signature(getattr(ref_qs, fn)).parameters)
                    kw = {k: v for k, v in extra.items() if k in allowed}
                    if extra and not kw:
                        continue
                    args = {"query_by_text_clip": ["person", 7], "query_by_keyword": ["Hotline", thr, 7],
                            "query_by_object": ["person", thr, 7]}.get(fn, ["person", thr, 7])
                    outs = []
                    for mod, ss, ds in ((qs, ours_ss, ours_ds), (ref_qs, ref_ss, ref_ds)):
                        cbs = {"search": search, "conf": conf, "kw": ss.search_frames_by_keyword,
                               "fmt": ds.format_event_for_frontend}
                        outs.append(getattr(mod, fn)(*args, *[cbs[c] for c in cb_names], video_name="v",
                                                     video_data_mapping=mapping, **kw))
                    assert outs[0] == outs[1], (fn, thr, kw, trial)
