"""Cross-checks oracle/vit_oracle.py (restatement of openai/CLIP clip/model.py VisionTransformer) against an
independent implementation that is installed offline: transformers.CLIPVisionModelWithProjection with
hidden_act="quick_gelu" (SURVEY.md 8c).  The OpenAI weights are not on this box, so the encode boundary is
"parity unpinned" by reference artefacts; this is the strongest available pin."""
import numpy as np
import pytest
import torch

from oracle import vit_oracle as vo


def _hf_from_openai(visual, name):
    tr = pytest.importorskip("transformers")
    g = vo.GEOMETRY[name]
    cfg = tr.CLIPVisionConfig(hidden_size=g["width"], intermediate_size=4 * g["width"],
                              num_hidden_layers=g["layers"], num_attention_heads=g["heads"],
                              image_size=g["image_resolution"], patch_size=g["patch_size"], hidden_act="quick_gelu",
                              projection_dim=g["embed_dim"], layer_norm_eps=1e-5, attention_dropout=0.0)
    hf = tr.CLIPVisionModelWithProjection(cfg).eval()
    sd = visual.state_dict()
    d = g["width"]
    new = {
        "vision_model.embeddings.class_embedding": sd["class_embedding"],
        "vision_model.embeddings.patch_embedding.weight": sd["conv1.weight"],
        "vision_model.embeddings.position_embedding.weight": sd["positional_embedding"],
        "vision_model.pre_layrnorm.weight": sd["ln_pre.weight"],
        "vision_model.pre_layrnorm.bias": sd["ln_pre.bias"],
        "vision_model.post_layernorm.weight": sd["ln_post.weight"],
        "vision_model.post_layernorm.bias": sd["ln_post.bias"],
        "visual_projection.weight": sd["proj"].t().contiguous(),
    }
    for i in range(g["layers"]):
        s, t = f"transformer.resblocks.{i}.", f"vision_model.encoder.layers.{i}."
        w, b = sd[s + "attn.in_proj_weight"], sd[s + "attn.in_proj_bias"]
        for j, nm in enumerate(("q_proj", "k_proj", "v_proj")):
            new[t + f"self_attn.{nm}.weight"] = w[j * d:(j + 1) * d]
            new[t + f"self_attn.{nm}.bias"] = b[j * d:(j + 1) * d]
        new[t + "self_attn.out_proj.weight"] = sd[s + "attn.out_proj.weight"]
        new[t + "self_attn.out_proj.bias"] = sd[s + "attn.out_proj.bias"]
        new[t + "layer_norm1.weight"], new[t + "layer_norm1.bias"] = sd[s + "ln_1.weight"], sd[s + "ln_1.bias"]
        new[t + "layer_norm2.weight"], new[t + "layer_norm2.bias"] = sd[s + "ln_2.weight"], sd[s + "ln_2.bias"]
        new[t + "mlp.fc1.weight"], new[t + "mlp.fc1.bias"] = sd[s + "mlp.c_fc.weight"], sd[s + "mlp.c_fc.bias"]
        new[t + "mlp.fc2.weight"], new[t + "mlp.fc2.bias"] = sd[s + "mlp.c_proj.weight"], sd[s + "mlp.c_proj.bias"]
    missing, unexpected = hf.load_state_dict(new, strict=False)
    missing = [m for m in missing if "position_ids" not in m]
    assert not missing and not unexpected, (missing, unexpected)
    return hf


def test_param_count_matches_openai_vit_b32():
    v = vo.build_visual("ViT-B/32", seed=1)
    n = sum(p.numel() for p in v.parameters())
    assert n == 87_849_216  # openai/CLIP ViT-B/32 visual tower
    keys = set(v.state_dict().keys())
    for k in ("conv1.weight", "class_embedding", "positional_embedding", "ln_pre.weight", "proj",
              "transformer.resblocks.11.attn.in_proj_weight", "transformer.resblocks.0.mlp.c_fc.bias"):
        assert k in keys
    assert v.state_dict()["positional_embedding"].shape == (50, 768)


def test_oracle_matches_hf_clip_vision():
    torch.manual_seed(0)
    name = "ViT-B/32"
    visual = vo.build_visual(name, seed=1)
    hf = _hf_from_openai(visual, name)
    _, pix = vo.synthetic_frames(3, seed=0)
    with torch.no_grad():
        ref = hf(pixel_values=pix).image_embeds.numpy()
    out = vo.encode_image(visual, pix, normalise=False)
    cos = (out * ref).sum(-1) / np.linalg.norm(out, axis=-1) / np.linalg.norm(ref, axis=-1)
    assert cos.min() > 0.99999
    assert np.abs(out - ref).max() < 1e-3 * max(1.0, np.abs(ref).max())


def test_synthetic_frames_are_deterministic_and_normalised():
    u8a, pa = vo.synthetic_frames(2, seed=0)
    u8b, pb = vo.synthetic_frames(2, seed=0)
    assert torch.equal(u8a, u8b) and torch.equal(pa, pb)
    assert u8a.shape == (2, 224, 224, 3) and pa.shape == (2, 3, 224, 224)
    x = u8a[0, 5, 7, 1].item() / 255.0
    assert abs(pa[0, 1, 5, 7].item() - (x - vo.CLIP_MEAN[1]) / vo.CLIP_STD[1]) < 1e-6


def test_encode_image_normalises_like_service():
    # embedding_service.py:499-502: vstack then / np.linalg.norm(axis=-1, keepdims=True)
    visual = vo.build_visual("ViT-B/32", seed=1)
    _, pix = vo.synthetic_frames(2, seed=3)
    e = vo.encode_image(visual, pix, normalise=True)
    assert e.shape == (2, 512)
    assert np.allclose(np.linalg.norm(e, axis=-1), 1.0, atol=1e-5)
