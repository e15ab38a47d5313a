"""Drop-in surface on the GPU: the reference's call sequence (ingest -> get_embeddings -> search_top_frames ->
extract_query_confidence) through `services.EmbeddingService` / `clip` / `embedding`, checked against the oracle.

Tolerances: encode cosine >= 0.999 per row vs the fp32 oracle; search ranks identical up to ties within 2^-10.
"""
import importlib
import json
import os

import numpy as np
import pytest
import torch

from conftest import PKG
from oracle import search_oracle as so
from oracle import vit_oracle as vo

pytestmark = pytest.mark.gpu
EPS = 2.0 ** -10


@pytest.fixture(scope="module")
def visual():
    return vo.build_visual("ViT-B/32", seed=1)


@pytest.fixture(scope="module")
def full_sd(visual):
    sd = {"visual." + k: v for k, v in visual.state_dict().items()}
    sd["logit_scale"] = torch.tensor(2.6592)
    return sd


def _write_frames(folder, names, seed=0, size=(256, 320)):
    from PIL import Image
    rng = np.random.default_rng(seed)
    os.makedirs(folder, exist_ok=True)
    for n in names:
        # smooth random images survive PNG exactly and resize without aliasing surprises
        low = rng.integers(0, 255, (8, 10, 3), dtype=np.uint8)
        Image.fromarray(low).resize((size[1], size[0]), Image.BICUBIC).save(os.path.join(folder, n))


def _site(tmp_path, full_sd, videos, text_encoder=None):
    services = importlib.import_module(PKG + ".services")
    meta, emb = tmp_path / "metadata", tmp_path / "embedding"
    meta.mkdir(exist_ok=True)
    emb.mkdir(exist_ok=True)
    mapping = {}
    for v, names in videos.items():
        frames_dir = tmp_path / "frames" / v
        _write_frames(str(frames_dir), names, seed=sum(map(ord, v)))
        items = [{"filepath": str(frames_dir / n), "frameidx": int(os.path.splitext(n)[0])} for n in sorted(names)]
        (meta / f"{v}_metadata.json").write_text(json.dumps(items))
        mapping[v] = {"metadata_file": str(meta / f"{v}_metadata.json"),
                      "embeddings_file": str(emb / f"{v}_embeddings.npy"),
                      "video_path": "", "frames_dir": str(frames_dir)}
    (meta / "video_mapping.json").write_text(json.dumps(mapping))
    ps = services.PathService(str(tmp_path / "frames"), str(meta), str(emb))
    cs = services.CacheService()
    ds = services.DataService(ps, cs)
    svc = services.EmbeddingService(cs, ps, ds, "cuda", load_kwargs=dict(state_dict=full_sd, max_batch=64),
                                    text_encoder=text_encoder)
    return svc, ps


def _oracle_rows(visual, svc, folder, names):
    from PIL import Image
    pix = torch.stack([svc.preprocess(Image.open(os.path.join(folder, n)).convert("RGB")) for n in names])
    return vo.encode_image(visual, pix, normalise=True)


def test_ingest_writes_reference_layout(tmp_path, cuda, visual, full_sd):
    names = [f"{i}.png" for i in (10074, 1061, 7, 300, 25, 42, 9000)] + ["notes.txt"]
    svc, ps = _site(tmp_path, full_sd, {"vid": [n for n in names if n.endswith(".png")]})
    folder = ps.video_data_mapping["vid"]["frames_dir"]
    open(os.path.join(folder, "notes.txt"), "w").write("not an image")
    path = svc.extract_and_save_embeddings_from_folder(folder, None, "vid")
    assert path == ps.get_embeddings_path("vid") and os.path.exists(path)
    out = np.load(path)
    order = sorted(n for n in names if n.endswith(".png"))  # lexicographic: 10074 < 1061 < 25 ...
    assert order[:2] == ["10074.png", "1061.png"]
    assert out.shape == (7, 512) and out.dtype in (np.float16, np.float32) and out.flags["C_CONTIGUOUS"]
    assert np.allclose(np.linalg.norm(out.astype(np.float32), axis=-1), 1.0, atol=2e-3)
    ref = _oracle_rows(visual, svc, folder, order)
    cos = (out.astype(np.float32) * ref).sum(-1) / np.linalg.norm(out.astype(np.float32), axis=-1)
    assert cos.min() >= 0.999, cos
    meta = json.load(open(ps.get_metadata_path("vid")))
    assert all(item["embedding_model"] == "original" for item in meta)
    assert svc.get_active_model_name() == "original"
    # the device resize/crop path (default) and the per-frame host `preprocess` path see identical pixels
    assert svc.gpu_preprocess
    svc.gpu_preprocess = False
    host = np.load(svc.extract_and_save_embeddings_from_folder(folder, None, "vid"))
    assert np.abs(host.astype(np.float32) - out.astype(np.float32)).max() <= 2e-3


def test_ingest_edge_cases(tmp_path, cuda, full_sd):
    svc, ps = _site(tmp_path, full_sd, {"vid": ["1.png", "2.png", "3.png"]})
    empty = tmp_path / "empty"
    empty.mkdir()
    assert svc.extract_and_save_embeddings_from_folder(str(empty), None, "vid") is None
    folder = ps.video_data_mapping["vid"]["frames_dir"]
    open(os.path.join(folder, "2.png"), "wb").write(b"broken")  # undecodable frame -> zero row, no NaN
    out = np.load(svc.extract_and_save_embeddings_from_folder(folder, "finetuned", "vid"))  # falls back to original
    assert np.isfinite(out).all() and np.all(out[1] == 0) and np.all(np.abs(out[[0, 2]]).sum(-1) > 0)
    # a video that is not in the mapping yet gets its own file instead of overwriting the default store (8g)
    new = np.load(svc.extract_and_save_embeddings_from_folder(folder, None, "fresh"))
    assert os.path.exists(os.path.join(ps.embedding_dir, "fresh_embeddings.npy")) and new.shape == (3, 512)
    assert svc.extract_image_embedding(os.path.join(folder, "2.png")) is None
    one = svc.extract_image_embedding(os.path.join(folder, "1.png"))
    assert one.shape == (1, 512) and abs(np.linalg.norm(one.astype(np.float32)) - 1) < 2e-3
    assert np.allclose(one.astype(np.float32)[0], out[0].astype(np.float32), atol=4e-3)


def test_search_calls_match_oracle(tmp_path, cuda, full_sd):
    rng = np.random.default_rng(3)
    table = {}

    def text_encoder(query):
        if query not in table:
            table[query] = rng.standard_normal((1, 512)).astype(np.float32)
        return table[query]

    names = [f"{i}.png" for i in range(3)]
    svc, ps = _site(tmp_path, full_sd, {"first": names, "second": names}, text_encoder)
    # the stores: un-normalised fp32 (first / default) and fp16 (second), 500 rows each, metadata to match
    stores = {}
    for v, dt in (("first", np.float32), ("second", np.float16)):
        raw = (rng.standard_normal((500, 512)) * 3).astype(dt)
        np.save(ps.get_embeddings_path(v), raw)
        items = [{"filepath": f"/x/{v}/{i}.jpg", "frameidx": i} for i in range(500)]
        json.dump(items, open(ps.get_metadata_path(v), "w"))
        stores[v] = raw
    for video, key in ((None, "first"), ("second", "second"), ("unknown", "first")):
        emb = svc.get_embeddings(video)
        ref_emb = so.normalise_store(stores[key])
        assert emb.shape == (500, 512) and emb.dtype == stores[key].dtype
        assert np.allclose(emb.astype(np.float32), ref_emb.astype(np.float32), atol=2 * EPS)
        for query in ("a person running", "xe máy"):
            q = svc.get_text_features(query, video)
            assert q.shape == (1, 512) and abs(np.linalg.norm(q) - 1) < 1e-5
            got = svc.search_top_frames(query, 30, video)
            pairs = svc.search_with_scores(q, 30, video)
            assert got == [f for f, _ in pairs] and len(got) == 30
            ref_i, ref_s = so.batched_topk(ref_emb.astype(np.float32), q, 30)
            got_i = np.array([int(os.path.splitext(f)[0]) for f in got])
            assert so.topk_equal_up_to_ties(got_i, np.array([s for _, s in pairs]), ref_i[0], ref_s[0], 2 * EPS)
            assert svc.search_top_frames(query, 30, video) is got          # memoised (:298, :339)
            assert svc.search_top_frames_by_image(q, 30, video) == got
            conf = svc.extract_query_confidence(got[0], query, video)
            assert isinstance(conf, float)
            assert abs(conf - so.query_confidence(ref_emb.astype(np.float32), int(got_i[0]), q)) < 2 * EPS
            assert abs(svc.extract_query_confidence("/some/dir/" + got[3], query, video) - pairs[3][1]) < 1e-5
            assert svc.extract_query_confidence("nope.jpg", query, video) == 0.0
    assert len(svc.search_top_frames("k larger than the store", 5000, None)) == 500


def test_errors_become_empty_results(tmp_path, cuda, full_sd):
    svc, ps = _site(tmp_path, full_sd, {"vid": ["1.png"]})   # no .npy written, no text tower, no text_encoder
    assert svc.get_embeddings("vid") is None
    assert svc.search_top_frames("anything", 10, "vid") == []
    assert svc.search_top_frames_by_image(np.ones((1, 512), np.float32), 10, "vid") == []
    assert svc.extract_query_confidence("1.png", "anything", "vid") == 0.0
    assert svc.set_active_model("finetuned") is False and svc.set_active_model("bogus") is False
    assert svc.set_active_model("original") is True


def test_clip_face_and_legacy_ingest(tmp_path, cuda, visual, full_sd):
    clip = importlib.import_module(PKG + ".clip")
    legacy = importlib.import_module(PKG + ".embedding")
    model, preprocess = clip.load("ViT-B/32", device="cuda", state_dict=full_sd, max_batch=8)
    assert model.visual.output_dim == 512 and model.dtype == torch.float16 and model.eval() is model
    _, pix = vo.synthetic_frames(20, seed=4)          # 20 > max_batch: encode_image chunks internally
    ref = vo.encode_image(visual, pix, normalise=False)
    out = model.encode_image(pix.to(cuda))
    assert out.dtype == torch.float16 and out.shape == (20, 512) and out.is_cuda
    out32 = model.float().encode_image(pix.to(cuda))
    assert out32.dtype == torch.float32
    o, r = out32.cpu().numpy(), ref
    cos = (o * r).sum(-1) / (np.linalg.norm(o, axis=-1) * np.linalg.norm(r, axis=-1))
    assert cos.min() >= 0.999
    assert np.allclose(np.linalg.norm(o, axis=-1), np.linalg.norm(r, axis=-1), rtol=2e-2)  # un-normalised, as clip
    with pytest.raises(NotImplementedError):
        model.encode_text(torch.zeros((1, 77), dtype=torch.long))
    # legacy entry point: un-normalised rows in the model dtype, `<video>_embeddings.npy`, os.walk order
    folder = tmp_path / "clip_7"
    _write_frames(str(folder), ["b.png", "a.jpg", "c.JPEG"], seed=9)
    (folder / "skip.txt").write_text("x")
    path = legacy.extract_and_save_embeddings_from_folder(
        str(folder), "ViT-B/32", embeddings_dir=str(tmp_path / "emb"), load_kwargs=dict(state_dict=full_sd))
    assert os.path.basename(path) == "clip_7_embeddings.npy"
    rows = np.load(path)
    assert rows.shape == (3, 512) and rows.dtype == np.float16
    assert np.linalg.norm(rows.astype(np.float32), axis=-1).min() > 2.0   # not normalised (embedding.py:55)


def test_text_query_strategies(tmp_path, cuda, full_sd):
    """query_strategies.query_by_text_clip / _with_adaptive_threshold over the service callbacks (a12): candidates =
    top_k*3 of the scan, joined to metadata by frameidx, scored by extract_query_confidence, best top_k returned."""
    qs = importlib.import_module(PKG + ".query_strategies")
    rng = np.random.default_rng(21)
    vec = rng.standard_normal((1, 512)).astype(np.float32)
    svc, ps = _site(tmp_path, full_sd, {"vid": ["0.png"]}, text_encoder=lambda q: vec)
    raw = rng.standard_normal((400, 512)).astype(np.float32)
    np.save(ps.get_embeddings_path("vid"), raw)
    items = [{"filepath": f"/x/vid/{i}.jpg", "frameidx": i, "tag": f"t{i}"} for i in range(400)]
    items.append({"filepath": "/x/vid/dup.jpg", "frameidx": 5, "tag": "second record with frameidx 5"})
    json.dump(items[:400], open(ps.get_metadata_path("vid"), "w"))
    fmt = lambda rec: {"id": rec["frameidx"], "title": rec["tag"], "confidence": rec["clip_similarity"]}
    emb = so.normalise_store(raw)
    q = vec / np.linalg.norm(vec)
    sim = (emb @ q.T).ravel()
    order = np.argsort(-sim)
    out = qs.query_by_text_clip("người đang chạy", 10, svc.search_top_frames, svc.extract_query_confidence, fmt, "vid",
                                ps.video_data_mapping)
    assert [e["id"] for e in out] == order[:10].tolist()
    assert all(abs(e["clip_similarity"] - sim[e["id"]]) < 2 * EPS and e["confidence"] == e["clip_similarity"] for e in out)
    assert all(a["clip_similarity"] >= b["clip_similarity"] for a, b in zip(out, out[1:]))
    thr = float(sim[order[4]]) - 1e-4                               # five frames pass the threshold
    out = qs.query_by_text_with_adaptive_threshold("q", thr, 10, svc.search_top_frames, svc.extract_query_confidence,
                                                   fmt, "vid", ps.video_data_mapping)
    assert [e["id"] for e in out] == order[:5].tolist()
    # unknown video + no default metadata file -> [] (errors never propagate)
    qs.METADATA_DIR = str(tmp_path / "nowhere")
    assert qs.query_by_text_clip("q", 5, svc.search_top_frames, svc.extract_query_confidence, fmt, "other", {}) == []
    qs.text_processor = lambda s: s.upper()
    seen = []
    qs.query_by_text_clip("abc", 2, lambda t, k, v: seen.append((t, k)) or [], svc.extract_query_confidence, fmt, "vid",
                          ps.video_data_mapping)
    qs.text_processor = None
    assert seen == [("ABC", 6)]                                     # processed text, top_k * 3 candidates


def test_search_by_image(tmp_path, cuda, visual, full_sd):
    """a15: query by image = encode -> scan(top_k*3) -> threshold -> top_k; a frame of the video finds itself first."""
    qs = importlib.import_module(PKG + ".query_strategies")
    names = [f"{i}.png" for i in range(12)]
    svc, ps = _site(tmp_path, full_sd, {"vid": names})
    folder = ps.video_data_mapping["vid"]["frames_dir"]
    assert svc.extract_and_save_embeddings_from_folder(folder, None, "vid") is not None
    fmt = lambda rec: {"id": rec["frameidx"]}
    query = os.path.join(folder, "7.png")
    out = qs.search_by_image(svc, query, -1.0, 3, fmt, "vid", ps.video_data_mapping)
    assert len(out) == 3 and out[0]["id"] == 7 and abs(out[0]["clip_similarity"] - 1.0) < 4e-3
    assert out[0]["confidence"] == out[0]["clip_similarity"]
    assert all(a["clip_similarity"] >= b["clip_similarity"] for a, b in zip(out, out[1:]))
    # oracle: cosines of the query frame's fp32 embedding against the stored rows, ranks up to near-ties
    order = sorted(names)
    ref = _oracle_rows(visual, svc, folder, order)
    sim = ref @ ref[order.index("7.png")]
    assert abs(out[1]["clip_similarity"] - np.sort(sim)[-2]) < 4e-3
    only_self = qs.search_by_image(svc, query, float(np.sort(sim)[-2]) + 0.5 * (1 - float(np.sort(sim)[-2])), 3, fmt, "vid",
                                   ps.video_data_mapping)
    assert [e["id"] for e in only_self] == [7]
    from PIL import Image
    assert qs.search_by_image(svc, Image.open(query), -1.0, 1, fmt, "vid", ps.video_data_mapping)[0]["id"] == 7
    assert qs.search_by_image(svc, "http://example.com/x.jpg", 0.0, 3, fmt, "vid", ps.video_data_mapping) == []
    assert qs.search_by_image(svc, str(tmp_path / "missing.jpg"), 0.0, 3, fmt, "vid", ps.video_data_mapping) == []


def test_finetuned_checkpoint_switch(tmp_path, cuda, full_sd):
    """`"model": "finetuned"` (embedding_service.py:93-120): a checkpoint written the way the reference's training
    script writes it -- {'model_state_dict': CLIPWithClassifier.state_dict()} = `clip_model.*` keys + the classifier
    head -- is loaded into the B200 tower; ingest with the fine-tuned model differs from the original one and equals
    the oracle run on the fine-tuned weights (forward = encode_image -> fp32 -> L2 normalise, :36-49)."""
    services = importlib.import_module(PKG + ".services")
    tuned = vo.build_visual("ViT-B/32", seed=77)             # "fine-tuned" vision weights: another seed
    ckpt = {"clip_model.visual." + k: v.clone() for k, v in tuned.state_dict().items()}
    ckpt["clip_model.logit_scale"] = torch.tensor(2.6592)
    ckpt["logit_scale"] = torch.tensor(2.6592)
    ckpt.update({"classifier.0.weight": torch.zeros(512, 512), "classifier.0.bias": torch.zeros(512),
                 "classifier.3.weight": torch.zeros(3, 512), "classifier.3.bias": torch.zeros(3)})
    path = str(tmp_path / "final_checkpoint.pt")
    torch.save({"model_state_dict": ckpt, "epoch": 3}, path)

    names = [f"{i}.png" for i in (3, 1, 2, 10)]
    svc, ps = _site(tmp_path, full_sd, {"vid": names})
    assert svc.set_active_model("finetuned") is False        # nothing loaded yet
    svc._load_finetuned_model(path)
    assert svc.set_active_model("finetuned") is True and svc.get_active_model_name() == "finetuned"
    folder = os.path.join(str(tmp_path), "frames", "vid")
    out_ft = np.load(svc.extract_and_save_embeddings_from_folder(folder, "finetuned", "vid"))
    out_or = np.load(svc.extract_and_save_embeddings_from_folder(folder, "original", "vid"))
    order = sorted(names)
    ref_ft = _oracle_rows(tuned, svc, folder, order)
    cos = (out_ft.astype(np.float32) * ref_ft).sum(-1)
    assert cos.min() >= 0.999, cos
    assert ((out_ft.astype(np.float32) * out_or.astype(np.float32)).sum(-1) < 0.9).all()  # really another model
    meta = json.load(open(ps.get_metadata_path("vid")))
    assert all(item["embedding_model"] == "original" for item in meta)  # stamped by the last ingest
    # constructor path: EmbeddingService(checkpoint_path=...) loads it at start-up like :96-101
    cs = services.CacheService()
    svc2 = services.EmbeddingService(cs, ps, services.DataService(ps, cs), "cuda",
                                     load_kwargs=dict(state_dict=full_sd, max_batch=64), checkpoint_path=path)
    assert svc2.finetuned_model is not None and svc2.set_active_model("finetuned") is True


def test_ingest_video_end_to_end(tmp_path, cuda, full_sd):
    """The processing half of /api/upload-video (app.py:636-659): video file -> scene-cut key-frames on the GPU ->
    embeddings -> metadata -> video_mapping.json; the new video is then searchable by one of its own key-frames."""
    cv2 = pytest.importorskip("cv2")
    from oracle import scenecut_oracle as sco
    services = importlib.import_module(PKG + ".services")
    clip = sco.synthetic_clip(75, h=96, w=128, cuts=(20, 41, 60), seed=12)
    video = str(tmp_path / "demo.avi")
    wr = cv2.VideoWriter(video, cv2.VideoWriter_fourcc(*"MJPG"), 25.0, (128, 96))
    if not wr.isOpened():
        pytest.skip("no MJPG writer in this cv2 build")
    for f in clip:
        wr.write(f)
    wr.release()
    svc, ps = _site(tmp_path, full_sd, {})
    entry = services.ingest_video(svc, video, threshold=20.0)
    assert set(entry) == {"metadata_file", "embeddings_file", "video_path", "frames_dir", "embedding_model"}
    frames = sorted(os.listdir(entry["frames_dir"]))
    assert frames == sorted(["10.jpg", "30.jpg", "50.jpg", "67.jpg"])       # middle frames of the 4 scenes
    emb = np.load(entry["embeddings_file"])
    assert emb.shape == (4, 512) and np.allclose(np.linalg.norm(emb, axis=-1), 1.0, atol=2e-3)
    meta = json.load(open(entry["metadata_file"]))
    assert [os.path.basename(m["filepath"]) for m in meta] == frames and meta[0]["frameidx"] == 10
    assert meta[0]["video"] == video and meta[0]["object_detections"] == {"detections": []}
    mapping = json.load(open(os.path.join(ps.metadata_dir, "video_mapping.json")))
    assert mapping["demo"]["embeddings_file"] == entry["embeddings_file"]
    # searchable: each key-frame finds itself first
    for row, name in enumerate(frames):
        top = svc.search_top_frames_by_image(emb[row:row + 1].astype(np.float32), 2, "demo")
        assert top[0] == name
    with pytest.raises(RuntimeError):   # a still clip has no cut -> no key-frame
        still = str(tmp_path / "still.avi")
        wr = cv2.VideoWriter(still, cv2.VideoWriter_fourcc(*"MJPG"), 25.0, (128, 96))
        for _ in range(20):
            wr.write(clip[0])
        wr.release()
        services.ingest_video(svc, still, threshold=20.0)
