"""Pure-PyTorch CPU fp32 restatement of openai/CLIP's vision tower.

TEST INFRASTRUCTURE (checker), never the shipped path.

The reference calls `model.encode_image` of the `clip` package
(/root/reference/Backend/embedding.py:49, services/embedding_service.py:41,414,490).
That package is the un-vendored submodule Backend/CLIP = github.com/openai/CLIP pinned at
dcba3cb2e2827b402d2701e7e1c7d9fed8a20ef1 (status "missing" in .SUBMODULES.json) and no
OpenAI weights are on this box => PARITY UNPINNED at the clip.encode_image boundary by
the reference's own artefacts.  This file restates the published algorithm of
clip/model.py (LayerNorm-in-fp32, QuickGELU, ResidualAttentionBlock, Transformer,
VisionTransformer, CLIP.encode_image) with openai-compatible state_dict keys and is
cross-checked against transformers.CLIPVisionModelWithProjection (tests/test_oracle_vit.py).
"""
import math
from collections import OrderedDict

import torch
import torch.nn as nn
import torch.nn.functional as F

CLIP_MEAN = (0.48145466, 0.4578275, 0.40821073)   # compare_models.py:39,226,390
CLIP_STD = (0.26862954, 0.26130258, 0.27577711)

GEOMETRY = {  # clip/model.py build_model
    "ViT-B/32": dict(embed_dim=512, image_resolution=224, layers=12, width=768, patch_size=32, heads=12),
    "ViT-B/16": dict(embed_dim=512, image_resolution=224, layers=12, width=768, patch_size=16, heads=12),
    "ViT-L/14": dict(embed_dim=768, image_resolution=224, layers=24, width=1024, patch_size=14, heads=16),
}


class LayerNorm(nn.LayerNorm):
    def forward(self, x):
        t = x.dtype
        return super().forward(x.type(torch.float32)).type(t)


class QuickGELU(nn.Module):
    def forward(self, x):
        return x * torch.sigmoid(1.702 * x)


class ResidualAttentionBlock(nn.Module):
    def __init__(self, d, h):
        super().__init__()
        self.attn = nn.MultiheadAttention(d, h)
        self.ln_1 = LayerNorm(d)
        self.mlp = nn.Sequential(OrderedDict([
            ("c_fc", nn.Linear(d, d * 4)), ("gelu", QuickGELU()), ("c_proj", nn.Linear(d * 4, d))]))
        self.ln_2 = LayerNorm(d)

    def forward(self, x):
        y = self.ln_1(x)
        x = x + self.attn(y, y, y, need_weights=False)[0]
        return x + self.mlp(self.ln_2(x))


class Transformer(nn.Module):
    def __init__(self, width, layers, heads):
        super().__init__()
        self.width, self.layers = width, layers
        self.resblocks = nn.Sequential(*[ResidualAttentionBlock(width, heads) for _ in range(layers)])

    def forward(self, x):
        return self.resblocks(x)


class VisionTransformer(nn.Module):
    def __init__(self, image_resolution, patch_size, width, layers, heads, embed_dim):
        super().__init__()
        self.input_resolution, self.output_dim = image_resolution, embed_dim
        self.patch_size, self.width, self.heads, self.layers = patch_size, width, heads, layers
        self.conv1 = nn.Conv2d(3, width, kernel_size=patch_size, stride=patch_size, bias=False)
        scale = width ** -0.5
        self.class_embedding = nn.Parameter(scale * torch.randn(width))
        self.positional_embedding = nn.Parameter(scale * torch.randn((image_resolution // patch_size) ** 2 + 1, width))
        self.ln_pre = LayerNorm(width)
        self.transformer = Transformer(width, layers, heads)
        self.ln_post = LayerNorm(width)
        self.proj = nn.Parameter(scale * torch.randn(width, embed_dim))

    def forward(self, x):
        x = self.conv1(x)                                   # [B, d, g, g]
        x = x.reshape(x.shape[0], x.shape[1], -1).permute(0, 2, 1)   # [B, g*g, d]
        cls = self.class_embedding.to(x.dtype) + torch.zeros(x.shape[0], 1, x.shape[-1], dtype=x.dtype)
        x = torch.cat([cls, x], dim=1) + self.positional_embedding.to(x.dtype)
        x = self.ln_pre(x)
        x = x.permute(1, 0, 2)                              # [L, B, d]
        x = self.transformer(x)
        x = x.permute(1, 0, 2)
        x = self.ln_post(x[:, 0, :])
        return x @ self.proj


def build_visual(name="ViT-B/32", seed=1):
    """Seeded random init with openai's init scales (clip/model.py initialize_parameters)."""
    g = GEOMETRY[name]
    gen = torch.Generator().manual_seed(seed)
    m = VisionTransformer(g["image_resolution"], g["patch_size"], g["width"], g["layers"], g["heads"], g["embed_dim"])
    d, L = g["width"], g["layers"]
    proj_std, attn_std, fc_std = (d ** -0.5) * ((2 * L) ** -0.5), d ** -0.5, (2 * d) ** -0.5
    with torch.no_grad():
        def rn(t, std):
            t.copy_(torch.randn(t.shape, generator=gen) * std)
        rn(m.conv1.weight, (3 * g["patch_size"] ** 2) ** -0.5)
        rn(m.class_embedding, d ** -0.5)
        rn(m.positional_embedding, d ** -0.5)
        rn(m.proj, d ** -0.5)
        for blk in m.transformer.resblocks:
            rn(blk.attn.in_proj_weight, attn_std)
            rn(blk.attn.in_proj_bias, 0.02)
            rn(blk.attn.out_proj.weight, proj_std)
            rn(blk.attn.out_proj.bias, 0.02)
            rn(blk.mlp.c_fc.weight, fc_std)
            rn(blk.mlp.c_fc.bias, 0.02)
            rn(blk.mlp.c_proj.weight, proj_std)
            rn(blk.mlp.c_proj.bias, 0.02)
            for ln in (blk.ln_1, blk.ln_2):
                ln.weight.copy_(1.0 + 0.1 * torch.randn(d, generator=gen))
                ln.bias.copy_(0.05 * torch.randn(d, generator=gen))
        for ln in (m.ln_pre, m.ln_post):
            ln.weight.copy_(1.0 + 0.1 * torch.randn(d, generator=gen))
            ln.bias.copy_(0.05 * torch.randn(d, generator=gen))
    return m.eval()


def synthetic_frames(n, seed=0, res=224):
    """uint8 RGB uniform -> /255 -> CLIP mean/std -> [n,3,res,res] fp32."""
    gen = torch.Generator().manual_seed(seed)
    u8 = torch.randint(0, 256, (n, res, res, 3), generator=gen, dtype=torch.uint8)
    return u8, normalise_u8(u8)


def normalise_u8(u8):
    x = u8.permute(0, 3, 1, 2).float() / 255.0
    mean = torch.tensor(CLIP_MEAN).view(1, 3, 1, 1)
    std = torch.tensor(CLIP_STD).view(1, 3, 1, 1)
    return (x - mean) / std


@torch.no_grad()
def encode_image(visual, pixels, batch=32, normalise=True):
    """Driven like embedding_service.py:467-502 (batches of 32, .cpu().numpy(), final L2 norm)."""
    import numpy as np
    outs = []
    for i in range(0, pixels.shape[0], batch):
        outs.append(visual(pixels[i:i + batch].float()).cpu().numpy())
    e = np.vstack(outs)
    if normalise:
        e = e / np.linalg.norm(e, axis=-1, keepdims=True)
    return e
