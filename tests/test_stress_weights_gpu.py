"""Towers under weights shaped like TRAINED CLIP checkpoints rather than fresh initialisations.

Seeded random-init weights (LayerNorm gain 1, no outliers) are the gentlest case for a bf16 tensor-core path.  Trained
CLIP towers carry a few residual channels 50-100x larger than the rest (fed by positional-embedding / c_proj-bias
outliers), LayerNorm gains far from 1 and real biases; the reference runs them in fp16 (original) or fp32 (fine-tuned
wrapper) while this path carries bf16 activations between fp32-accumulated GEMMs with an fp32 residual stream.  No
trained weights are on the box (no network), so the regime is synthesised: the oracle's seeded towers are perturbed
in place, and the same state_dict goes to the fp32 oracle and to the GPU.

Asserted: cosine >= 0.999 per row against the fp32 oracle (north_star), and ranking preserved through the fused scan --
R@1 of noisy-copy queries identical, and every frame of the GPU store's top-10 scores within the embedding error of the
oracle store's 10th best (top-k identical up to ties within the stated tolerance).  Real-weight parity itself stays open; tests/test_real_weights_gpu.py closes it
wherever the OpenAI file exists.
"""
import numpy as np
import pytest
import torch

from oracle import search_oracle as so
from oracle import text_oracle as to
from oracle import vit_oracle as vo

pytestmark = pytest.mark.gpu
MIN_COS = 0.999


def _cos(a, b):
    return (a * b).sum(-1) / np.linalg.norm(a, axis=-1) / np.linalg.norm(b, axis=-1)


def stress_(blocks, lns, pos, width, seed, typical):
    """In-place: outlier residual channels (`typical` = the stream's ordinary per-channel magnitude), LayerNorm gains /
    shifts far from (1, 0), biases of real size."""
    g = torch.Generator().manual_seed(seed)
    outliers = torch.randperm(width, generator=g)[:4]
    with torch.no_grad():
        pos[:, outliers] += torch.tensor([60.0, -80.0, 100.0, 50.0]) * typical       # 50-100 x the channel scale
        for ln in lns:
            ln.weight.copy_(torch.exp(0.6 * torch.randn(width, generator=g)))        # gains ~ 0.3 .. 3.5
            ln.weight[outliers] = torch.tensor([0.05, 0.1, 0.02, 0.2])                # trained towers damp them
            ln.bias.copy_(0.5 * torch.randn(width, generator=g))
        for blk in blocks:
            blk.mlp.c_proj.bias[outliers] += torch.tensor([3.0, -4.0, 5.0, 2.5]) * typical   # the outliers grow
            blk.attn.in_proj_bias.add_(0.3 * torch.randn(blk.attn.in_proj_bias.shape, generator=g))
            blk.mlp.c_fc.bias.add_(0.3 * torch.randn(blk.mlp.c_fc.bias.shape, generator=g))
            blk.attn.out_proj.bias.add_(0.1 * torch.randn(width, generator=g))
    return outliers


def _ranking_preserved(pkg, cuda, emb_gpu, emb_ref, min_cos, seed):
    """Top-10 through the fused scan over the GPU-encoded store vs the oracle's top-10 over the oracle store: the best
    hit is the same frame, and every frame the GPU list holds is within `tol` of making the oracle's list (ties within
    the embedding error: a row with cosine c to its fp32 twin moves a score by at most sqrt(2 (1 - c)))."""
    rng = np.random.default_rng(seed)
    n, d = emb_ref.shape
    picks = rng.choice(n, 24, replace=False)
    queries = emb_ref[picks] + 0.05 * rng.standard_normal((24, d)).astype(np.float32) / np.sqrt(d)
    queries /= np.linalg.norm(queries, axis=-1, keepdims=True)
    store = pkg.search.FrameStore.from_tensor(emb_gpu, device=cuda)
    sc, idx = store.search(torch.from_numpy(queries).to(cuda), 10)
    sims = so.normalise_store(emb_ref) @ queries.T                      # [n, 24] oracle scores
    ref_i, ref_s = so.batched_topk(so.normalise_store(emb_ref), queries, 10)
    idx = idx.cpu().numpy()
    tol = 2.0 ** -10 + float(np.sqrt(2 * (1 - min(min_cos, 1.0))))
    assert np.array_equal(idx[:, 0], ref_i[:, 0]) and np.array_equal(idx[:, 0], picks)   # R@1 unchanged
    for j in range(24):
        assert np.all(sims[idx[j], j] >= ref_s[j, -1] - tol), (j, sims[idx[j], j], ref_s[j])
    overlap = [len(set(idx[j]) & set(ref_i[j])) for j in range(24)]
    print(f"top-10 overlap with the oracle lists: min {min(overlap)}, mean {np.mean(overlap):.2f} (tol {tol:.1e})")


@pytest.mark.parametrize("name,n,batch", [("ViT-B/32", 96, 32), ("ViT-L/14", 40, 8)])
def test_vision_tower_with_outlier_channels(pkg, cuda, name, n, batch):
    visual = vo.build_visual(name, seed=11)
    blocks = list(visual.transformer.resblocks)
    lns = [visual.ln_pre, visual.ln_post] + [ln for b in blocks for ln in (b.ln_1, b.ln_2)]
    u8, pix = vo.synthetic_frames(n, seed=8)
    # frames with structure (smooth + noise) so embeddings differ by more than the tower's rounding
    low = torch.nn.functional.interpolate(pix[:, :, ::28, ::28], size=224, mode="bilinear")
    pix = 0.7 * low + 0.3 * pix

    def stream(p):   # the residual stream entering ln_pre
        with torch.no_grad():
            x = visual.conv1(p).flatten(2).permute(0, 2, 1)
            return torch.cat([visual.class_embedding + torch.zeros(p.shape[0], 1, visual.width), x], 1) \
                + visual.positional_embedding

    typical = float(stream(pix[:2]).abs().median())
    stress_(blocks, lns, visual.positional_embedding, visual.width, seed=21, typical=typical)
    x = stream(pix[:2])
    ratio = (x.abs().amax(-1) / x.abs().median(-1).values).min()
    assert ratio >= 50, float(ratio)   # the stream really is in the stressed regime: outliers >= 50 x the median
    ref = vo.encode_image(visual, pix, batch=batch, normalise=True)
    enc = pkg.vit.VisionEncoder.from_state_dict(visual.state_dict(), name, device=cuda, max_batch=batch)
    out = torch.cat([enc.encode(pix[i:i + batch].to(cuda), normalise=True) for i in range(0, n, batch)])
    cos = _cos(out.cpu().numpy(), ref)
    print(f"{name} stressed weights: min cosine {cos.min():.6f}, mean {cos.mean():.6f}")
    assert cos.min() >= MIN_COS, cos.min()
    _ranking_preserved(pkg, cuda, out, ref, float(cos.min()), seed=31)


def test_text_tower_with_outlier_channels(pkg, cuda):
    tower = to.build_text("ViT-B/32", seed=13)
    blocks = list(tower.transformer.resblocks)
    lns = [tower.ln_final] + [ln for b in blocks for ln in (b.ln_1, b.ln_2)]
    tokens = to.synthetic_tokens(64, seed=5)
    with torch.no_grad():
        typical = float((tower.token_embedding(tokens[:4]) + tower.positional_embedding).abs().median())
    stress_(blocks, lns, tower.positional_embedding, tower.positional_embedding.shape[1], seed=23, typical=typical)
    with torch.no_grad():
        x = tower.token_embedding(tokens[:4]) + tower.positional_embedding
    assert (x.abs().amax(-1) / x.abs().median(-1).values).min() >= 50
    ref = to.encode_text(tower, tokens, normalise=True)
    enc = pkg.text.TextEncoder.from_state_dict(tower.state_dict(), device=cuda, max_batch=16)
    out = torch.cat([enc.encode(tokens[i:i + 16].to(cuda), normalise=True) for i in range(0, 64, 16)])
    cos = _cos(out.cpu().numpy(), ref)
    print(f"text tower stressed weights: min cosine {cos.min():.6f}")
    assert cos.min() >= MIN_COS, cos.min()
    _ranking_preserved(pkg, cuda, out, ref, float(cos.min()), seed=33)
