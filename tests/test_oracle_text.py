"""Cross-checks oracle/text_oracle.py (restatement of openai/CLIP `CLIP.encode_text`) against an independent
implementation installed offline: transformers.CLIPTextModelWithProjection(hidden_act="quick_gelu").  The OpenAI
weights are absent, so the boundary is "parity unpinned" by reference artefacts; this is the strongest pin here."""
import numpy as np
import pytest
import torch

from oracle import text_oracle as to


def _hf_from_openai(tower, name):
    tr = pytest.importorskip("transformers")
    g = to.TEXT_GEOMETRY[name]
    cfg = tr.CLIPTextConfig(vocab_size=g["vocab_size"], hidden_size=g["width"], intermediate_size=4 * g["width"],
                            num_hidden_layers=g["layers"], num_attention_heads=g["heads"],
                            max_position_embeddings=g["context_length"], hidden_act="quick_gelu",
                            projection_dim=g["embed_dim"], layer_norm_eps=1e-5, attention_dropout=0.0,
                            eos_token_id=to.EOT, bos_token_id=to.SOT, pad_token_id=0)
    hf = tr.CLIPTextModelWithProjection(cfg).eval()
    sd = tower.state_dict()
    d = g["width"]
    new = {
        "text_model.embeddings.token_embedding.weight": sd["token_embedding.weight"],
        "text_model.embeddings.position_embedding.weight": sd["positional_embedding"],
        "text_model.final_layer_norm.weight": sd["ln_final.weight"],
        "text_model.final_layer_norm.bias": sd["ln_final.bias"],
        "text_projection.weight": sd["text_projection"].t().contiguous(),
    }
    for i in range(g["layers"]):
        s, t = f"transformer.resblocks.{i}.", f"text_model.encoder.layers.{i}."
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


def test_param_count_matches_openai_text_tower():
    tower = to.build_text("ViT-B/32")
    # CLIP ViT-B/32 has 151 277 313 parameters: 87 849 216 visual + 1 logit_scale + the text half
    assert sum(p.numel() for p in tower.parameters()) == 151_277_313 - 87_849_216 - 1


def test_oracle_matches_hf_clip_text():
    tower = to.build_text("ViT-B/32", seed=5)
    hf = _hf_from_openai(tower, "ViT-B/32")
    tokens = to.synthetic_tokens(6, seed=1)
    ours = to.encode_text(tower, tokens, normalise=False)
    with torch.no_grad():
        theirs = hf(input_ids=tokens, attention_mask=torch.ones_like(tokens)).text_embeds.numpy()
    assert np.abs(ours - theirs).max() < 1e-3
    cos = (ours * theirs).sum(-1) / np.linalg.norm(ours, axis=-1) / np.linalg.norm(theirs, axis=-1)
    assert cos.min() > 0.99999


def test_causal_mask_and_eot_row():
    tower = to.build_text("ViT-B/32", seed=5)
    tokens = to.synthetic_tokens(3, seed=2)
    base = to.encode_text(tower, tokens, normalise=False)
    eot = tokens.argmax(-1)
    changed = tokens.clone()
    for i in range(1, 3):                      # ids after the EOT position cannot influence the EOT row
        changed[i, eot[i] + 1:] = 17
    assert np.allclose(to.encode_text(tower, changed, normalise=False)[1:], base[1:], atol=1e-6)
    changed = tokens.clone()
    changed[:, 1] = (changed[:, 1] + 1) % 40000 + 1   # an id before it does
    assert np.abs(to.encode_text(tower, changed, normalise=False) - base).max() > 1e-4
    assert (tokens[0] != 0).all() and tokens[0, -1] == to.EOT   # row 0 uses the whole context
