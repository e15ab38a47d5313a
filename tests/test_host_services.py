"""Host-side service mirrors: path / cache / data rules that define result identity.

Runs on CPU.  Where /root/reference is present (this container) the mirrors are also cross-checked against the
reference's own stdlib-only PathService / CacheService classes imported unmodified.
"""
import importlib
import importlib.util
import json
import os
import time

import pytest

from conftest import PKG

REF = "/root/reference/Backend/services"


@pytest.fixture(scope="module")
def services():
    return importlib.import_module(PKG + ".services")


def _layout(tmp_path, videos=("video_b", "video_a")):
    meta, emb = tmp_path / "metadata", tmp_path / "embedding"
    meta.mkdir()
    emb.mkdir()
    mapping = {}
    for v in videos:
        frames = [{"filepath": f"/frames/{v}/{n}.jpg", "frameidx": n} for n in (10074, 1061, 7)]
        (meta / f"{v}_metadata.json").write_text(json.dumps(frames))
        mapping[v] = {"metadata_file": str(meta / f"{v}_metadata.json"),
                      "embeddings_file": str(emb / f"{v}_embeddings.npy"), "video_path": "", "frames_dir": ""}
    (meta / "video_mapping.json").write_text(json.dumps(mapping))
    return str(tmp_path), str(meta), str(emb)


def _ref_class(module, cls):
    path = os.path.join(REF, module + ".py")
    if not os.path.exists(path):
        pytest.skip("reference tree not present on this machine")
    spec = importlib.util.spec_from_file_location("ref_" + module, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return getattr(mod, cls)


def test_default_store_is_first_mapping_entry(services, tmp_path):
    base, meta, emb = _layout(tmp_path)
    ps = services.PathService(base, meta, emb)
    assert ps.get_embeddings_path(None).endswith("video_b_embeddings.npy")   # first key, not alphabetical
    assert ps.get_embeddings_path("nope").endswith("video_b_embeddings.npy")  # unknown video -> default
    assert ps.get_embeddings_path("video_a").endswith("video_a_embeddings.npy")
    assert ps.get_metadata_path("video_a").endswith("video_a_metadata.json")


def test_empty_mapping_falls_back_to_legacy_names(services, tmp_path):
    ps = services.PathService(str(tmp_path), str(tmp_path), str(tmp_path))
    assert os.path.basename(ps.get_embeddings_path()) == "image_embeddings.npy"
    assert os.path.basename(ps.get_metadata_path()) == "output_samples.json"


def test_path_service_matches_reference(services, tmp_path, capsys):
    Ref = _ref_class("path_service", "PathService")
    base, meta, emb = _layout(tmp_path)
    ours, ref = services.PathService(base, meta, emb), Ref(base, meta, emb)
    for v in (None, "", "video_a", "video_b", "missing"):
        assert ours.get_embeddings_path(v) == ref.get_embeddings_path(v)
        assert ours.get_metadata_path(v) == ref.get_metadata_path(v)
    ours.add_video_mapping("new", "m.json", "e.npy", "v.mp4", "frames")
    ref.add_video_mapping("new", "m.json", "e.npy", "v.mp4", "frames")
    assert ours.video_data_mapping == ref.video_data_mapping


def test_frames_follow_metadata_order(services, tmp_path):
    base, meta, emb = _layout(tmp_path)
    ps, cs = services.PathService(base, meta, emb), services.CacheService()
    ds = services.DataService(ps, cs)
    assert ds.load_frames_from_json("video_a") == ["10074.jpg", "1061.jpg", "7.jpg"]  # file order, basenames
    assert cs.get_frames_list(os.path.normpath(ps.get_metadata_path("video_a"))) is not None
    assert ds.load_frames_from_json("missing") == ds.load_frames_from_json(None)
    ps.video_data_mapping["broken"] = {"metadata_file": str(tmp_path / "absent.json"), "embeddings_file": "x"}
    assert ds.load_frames_from_json("broken") == []


def test_cache_keys_and_ttl(services):
    cs = services.CacheService(default_ttl=3600)
    cs.set_search_results("search_original_dog_30", None, ["a.jpg"])
    assert cs.get_search_results("search_original_dog_30", None) == ["a.jpg"]
    assert cs.get_search_results("search_original_dog_30", "v") is None       # video scopes the key
    assert cs.get_text_features("search_original_dog_30", None) is None       # caches are separate
    cs.set_text_features("q", "v", 1, ttl=0.05)
    assert cs.get_text_features("q", "v") == 1
    time.sleep(0.08)
    assert cs.get_text_features("q", "v") is None                             # expired entries vanish
    cs.set_path_cache("v", "e.npy", "m.json")
    assert cs.get_path_cache("v") == {"embeddings_path": "e.npy", "json_path": "m.json"}
    cs.clear_all_caches()
    assert cs.get_path_cache("v") is None


def test_cache_service_matches_reference(services):
    Ref = _ref_class("cache_service", "CacheService")
    ours, ref = services.CacheService(), Ref()
    for c in (ours, ref):
        c.set_search_results("k", "v", [1])
        c.set_text_features("k", None, 2)
        c.set_embeddings("p", 3)
        c.set_frames_list("j", [4])
    for probe in (lambda c: c.get_search_results("k", "v"), lambda c: c.get_search_results("k", None),
                  lambda c: c.get_text_features("k", None), lambda c: c.get_text_features("k", "default"),
                  lambda c: c.get_embeddings("p"), lambda c: c.get_frames_list("j"),
                  lambda c: c.get_frames_list("p"), lambda c: c.get_path_cache("v")):
        assert probe(ours) == probe(ref)


def test_clip_face_fails_loudly_without_weights_or_gpu(tmp_path):
    clip = importlib.import_module(PKG + ".clip")
    assert clip.available_models() == ["ViT-B/32", "ViT-B/16", "ViT-L/14"]
    with pytest.raises(RuntimeError, match="not found"):
        clip.load("RN50", device="cuda")
    os.environ.pop("EVR_CLIP_RANDOM_INIT", None)
    with pytest.raises(RuntimeError, match="no weights"):
        clip.load("ViT-B/32", device="cuda", download_root=str(tmp_path))
    with pytest.raises(RuntimeError, match="CUDA device only"):
        clip.load("ViT-B/32", device="cpu", random_init=1)
    sd = clip.random_visual_state_dict("ViT-B/32", seed=1)
    assert sum(v.numel() for k, v in sd.items() if k.startswith("visual.")) == 87_849_216


def test_threaded_decode_is_ordered_bounded_and_identical(services, tmp_path):
    """Ingest decodes on host threads (services.decode_frames): same pixels as the reference's serial
    `Image.open(p).convert('RGB')` (embedding_service.py:475), yielded in path order, a bad file comes back as the
    exception (-> zero row), and no more than `window` frames are ever decoded ahead of the consumer."""
    import numpy as np
    from PIL import Image
    rng = np.random.default_rng(0)
    paths = []
    for i in range(23):
        p = tmp_path / f"{i:04d}.jpg"
        Image.fromarray(rng.integers(0, 256, (40 + i, 64, 3), dtype=np.uint8)).save(p, quality=90)
        paths.append(str(p))
    bad = tmp_path / "0007.jpg"
    bad.write_bytes(b"not a jpeg")
    serial = list(services.decode_frames(paths, workers=1))
    threaded = list(services.decode_frames(paths, workers=4, window=5))
    assert len(serial) == len(threaded) == 23
    for i, (a, b) in enumerate(zip(serial, threaded)):
        if i == 7:
            assert isinstance(a, Exception) and isinstance(b, Exception)
        else:
            assert a.shape == (40 + i, 64, 3) and np.array_equal(a, b)
            assert np.array_equal(a, np.array(Image.open(paths[i]).convert('RGB')))
    # bounded lookahead: after consuming one frame at most `window` more have been opened
    opened = []
    real_open = Image.open

    def spy(path, *a, **k):
        opened.append(path)
        return real_open(path, *a, **k)
    Image.open = spy
    try:
        gen = services.decode_frames(paths, workers=2, window=3)
        next(gen)
        time.sleep(0.3)
        assert len(opened) <= 1 + 3
        gen.close()
    finally:
        Image.open = real_open


def test_frame_metadata_items_layout(services, tmp_path):
    """Stand-in for process_images_in_folder (JSON_sample_DOC.py:74-84, :90): same keys, sorted frame order (= the row
    order of the embedding file), frameidx from the digits of the file name."""
    from PIL import Image
    folder = tmp_path / "frames"
    folder.mkdir()
    for name in ("300.jpg", "7.jpg", "1061.jpg", "cover.png", "notes.txt"):
        if name.endswith(".txt"):
            (folder / name).write_text("x")
        else:
            Image.new("RGB", (32, 18), (10, 20, 30)).save(folder / name)
    out = services.frame_metadata_items(str(folder), str(tmp_path / "meta" / "v_metadata.json"), "/videos/v.mp4")
    items = json.load(open(out))
    assert [i["frameid"] for i in items] == sorted(["300.jpg", "7.jpg", "1061.jpg", "cover.png"])
    first = items[0]
    assert set(first) == {"id", "media_type", "filepath", "tags", "metadata", "video", "frameid", "text_detections",
                          "object_detections", "frameidx"}
    assert first["metadata"]["width"] == 32 and first["metadata"]["height"] == 18 and first["video"] == "/videos/v.mp4"
    assert set(first["metadata"]) == {"size_bytes", "mime_type", "width", "height", "num_channels"}   # :15-21
    assert {i["frameid"]: i["frameidx"] for i in items} == {"1061.jpg": 1061, "300.jpg": 300, "7.jpg": 7,
                                                           "cover.png": None}


def test_metadata_items_equal_the_references(services, tmp_path):
    """`frame_metadata_items` against tests/golden/ref_metadata_golden.json: the items the reference's own
    `process_image` / `get_image_metadata` (JSON_sample_DOC.py:12-87, run unmodified with detectors that find nothing)
    wrote for its own key-frames -- same keys in the same order, same values, `id` apart (a fresh uuid4 each)."""
    import shutil
    import uuid
    golden_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    with open(os.path.join(golden_dir, "ref_metadata_golden.json"), encoding="utf-8") as f:
        want = json.load(f)["items"]
    frames = tmp_path / "Violence"
    frames.mkdir()
    for name in os.listdir(os.path.join(golden_dir, "ref_frames")):
        if name.endswith(".jpg"):
            shutil.copyfile(os.path.join(golden_dir, "ref_frames", name), frames / name)
    (frames / "7.jpg").write_bytes((frames / want[0]["frameid"]).read_bytes())     # a numeric stem: frameidx = 7
    (frames / "notes.txt").write_text("not a frame")
    out = services.frame_metadata_items(str(frames), str(tmp_path / "meta" / "Violence_metadata.json"),
                                        "/videos/Violence.mp4")
    with open(out, encoding="utf-8") as f:
        got = json.load(f)
    assert [g["frameid"] for g in got] == sorted([w["frameid"] for w in want] + ["7.jpg"])
    by_name = {g["frameid"]: g for g in got}
    for w in want:
        g = by_name[w["frameid"]]
        assert list(g) == list(w)                                   # key order as the reference writes it
        assert uuid.UUID(g["id"]).version == 4
        assert os.path.basename(g["filepath"]) == w["filepath"] and os.path.isabs(g["filepath"])
        for key in w:
            if key not in ("id", "filepath"):
                assert g[key] == w[key], (w["frameid"], key)
    assert by_name["7.jpg"]["frameidx"] == 7 and by_name[want[0]["frameid"]]["frameidx"] is None
