"""Text tower on the GPU (`TextEncoder` / `model.encode_text`) against the fp32 oracle restatement of
openai/CLIP `CLIP.encode_text`; tolerance as for the vision tower: cosine >= 0.999 per row (bf16 compute)."""
import gzip
import importlib

import numpy as np
import pytest
import torch

from conftest import PKG
from oracle import text_oracle as to
from oracle import vit_oracle as vo

pytestmark = pytest.mark.gpu
MIN_COS = 0.999


def _cos(a, b):
    return (a * b).sum(-1) / np.linalg.norm(a, axis=-1) / np.linalg.norm(b, axis=-1)


@pytest.fixture(scope="module")
def tower():
    return to.build_text("ViT-B/32", seed=5)


def test_text_tower_matches_oracle(pkg, cuda, tower):
    tokens = to.synthetic_tokens(21, seed=3)
    ref = to.encode_text(tower, tokens, normalise=False)
    enc = pkg.text.TextEncoder.from_state_dict(tower.state_dict(), device=cuda, max_batch=8)  # chunks of 8, 8, 5
    out = enc.encode(tokens.to(cuda)).cpu().numpy()
    assert out.shape == (21, 512)
    cos = _cos(out, ref)
    assert cos.min() >= MIN_COS, cos
    assert np.abs(out - ref).max() <= 0.03 * np.abs(ref).max()
    # batch invariance, int32 ids, normalised fp16 output
    one = enc.encode(tokens[4:5].to(cuda).to(torch.int32))
    assert torch.equal(one[0], enc.encode(tokens.to(cuda))[4])
    o16 = enc.encode(tokens.to(cuda), normalise=True, out_dtype=torch.float16)
    assert o16.dtype == torch.float16
    assert np.allclose(np.linalg.norm(o16.float().cpu().numpy(), axis=-1), 1.0, atol=2e-3)


def test_causal_mask_on_device(pkg, cuda, tower):
    enc = pkg.text.TextEncoder.from_state_dict(tower.state_dict(), device=cuda, max_batch=4)
    tokens = to.synthetic_tokens(4, seed=9)
    base = enc.encode(tokens.to(cuda))
    changed = tokens.clone()
    eot = tokens.argmax(-1)
    for i in range(1, 4):
        changed[i, eot[i] + 1:] = 23            # after the EOT row: invisible under the causal mask
    assert torch.equal(enc.encode(changed.to(cuda))[1:], base[1:])
    changed = tokens.clone()
    changed[:, 1] = (changed[:, 1] + 5) % 40000 + 1
    assert (enc.encode(changed.to(cuda)) - base).abs().max() > 1e-4


def test_text_argument_errors(pkg, cuda, tower):
    enc = pkg.text.TextEncoder.from_state_dict(tower.state_dict(), device=cuda, max_batch=2)
    with pytest.raises(ValueError):
        enc.encode(torch.zeros((1, 76), dtype=torch.long, device=cuda))
    with pytest.raises(ValueError):
        enc.encode(torch.full((1, 77), 49408, dtype=torch.long, device=cuda))
    with pytest.raises(ValueError):
        enc.encode(torch.zeros((1, 77), device=cuda))
    assert enc.encode(torch.zeros((0, 77), dtype=torch.long, device=cuda)).shape == (0, 512)
    with pytest.raises(KeyError):
        pkg.text.TextEncoder.from_state_dict({"token_embedding.weight": torch.zeros(8, 512),
                                              "positional_embedding": torch.zeros(77, 512),
                                              "text_projection": torch.zeros(512, 512)}, device=cuda)


def test_clip_model_text_and_service_query_path(pkg, cuda, tower, tmp_path, monkeypatch):
    """`clip.load` with a full CLIP state_dict: encode_text in the model dtype, and the reference's query path
    (get_text_features -> search_top_frames -> extract_query_confidence) with the real tokenizer code on a
    synthetic merge table."""
    import json
    from oracle import search_oracle as so
    clip = importlib.import_module(PKG + ".clip")
    services = importlib.import_module(PKG + ".services")
    visual = vo.build_visual("ViT-B/32", seed=1)
    sd = {"visual." + k: v for k, v in visual.state_dict().items()}
    sd.update(tower.state_dict())
    model, _ = clip.load("ViT-B/32", device="cuda", state_dict=sd, max_batch=4)
    assert model.has_text_tower
    tokens = to.synthetic_tokens(3, seed=11)
    ref = to.encode_text(tower, tokens, normalise=False)
    out = model.encode_text(tokens.to(cuda))
    assert out.dtype == torch.float16 and _cos(out.float().cpu().numpy(), ref).min() >= MIN_COS
    # ---- service path
    merges = ["#version: synthetic", "r u", "ru n", "d o", "do g</w>"]
    with gzip.open(tmp_path / "bpe_simple_vocab_16e6.txt.gz", "wb") as f:
        f.write("\n".join(merges).encode())
    monkeypatch.setenv("EVR_CLIP_ROOT", str(tmp_path))
    meta, emb = tmp_path / "metadata", tmp_path / "embedding"
    meta.mkdir()
    emb.mkdir()
    rng = np.random.default_rng(4)
    raw = rng.standard_normal((300, 512)).astype(np.float32)
    np.save(emb / "v_embeddings.npy", raw)
    json.dump([{"filepath": f"/f/{i}.jpg", "frameidx": i} for i in range(300)], open(meta / "v_metadata.json", "w"))
    json.dump({"v": {"metadata_file": str(meta / "v_metadata.json"), "embeddings_file": str(emb / "v_embeddings.npy"),
                     "video_path": "", "frames_dir": ""}}, open(meta / "video_mapping.json", "w"))
    ps = services.PathService(str(tmp_path), str(meta), str(emb))
    cs = services.CacheService()
    svc = services.EmbeddingService(cs, ps, services.DataService(ps, cs), "cuda",
                                    load_kwargs=dict(state_dict=sd, max_batch=4))
    query = "a dog running"
    q = svc.get_text_features(query)
    tok = clip.tokenize([query])
    q_ref = to.encode_text(tower, tok, normalise=True)
    assert q.shape == (1, 512) and _cos(q.astype(np.float32), q_ref).min() >= MIN_COS
    got = svc.search_top_frames(query, 10)
    ref_i, ref_s = so.batched_topk(so.normalise_store(raw), q.astype(np.float32), 10)
    pairs = svc.search_with_scores(q, 10)
    got_i = np.array([int(f.split(".")[0]) for f in got])
    assert so.topk_equal_up_to_ties(got_i, np.array([s for _, s in pairs]), ref_i[0], ref_s[0], 2.0 ** -10)
    assert abs(svc.extract_query_confidence(got[0], query) - pairs[0][1]) < 1e-5


def test_single_query_graph_replay_is_identical(pkg, cuda, tower):
    """One query (the serving case) runs as a captured graph from its third call: same bits, new ids honoured."""
    enc = pkg.text.TextEncoder.from_state_dict(tower.state_dict(), device=cuda, max_batch=8)
    tokens = to.synthetic_tokens(6, seed=11).to(cuda)
    plain = [enc.encode(tokens[i:i + 1]).clone() for i in range(3)]          # call 1 plain, 2 capture, 3 replay
    again = [enc.encode(tokens[i:i + 1]).clone() for i in range(3)]          # all replays
    for p, a in zip(plain, again):
        assert torch.equal(p, a)
    whole = enc.encode(tokens)
    assert torch.equal(whole[:3], torch.cat(again))
    ref = to.encode_text(tower, tokens.cpu(), normalise=False)
    assert _cos(whole.cpu().numpy(), ref).min() >= MIN_COS
