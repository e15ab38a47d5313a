"""Pure-PyTorch CPU fp32 restatement of openai/CLIP's text tower (`CLIP.encode_text`).

TEST INFRASTRUCTURE (checker), never the shipped path.

The reference calls `model.encode_text(clip.tokenize([query]))`
(/root/reference/Backend/services/embedding_service.py:169-177) of the un-vendored `clip` package
(openai/CLIP @ dcba3cb2e2827b402d2701e7e1c7d9fed8a20ef1, absent; no OpenAI weights on the box) => PARITY UNPINNED by
reference artefacts at this boundary.  This file restates the published algorithm of clip/model.py (CLIP.__init__ text
half, build_attention_mask, encode_text) with openai state_dict keys and is cross-checked against
transformers.CLIPTextModelWithProjection(hidden_act="quick_gelu") in tests/test_oracle_text.py.
"""
from collections import OrderedDict

import numpy as np
import torch
import torch.nn as nn

from .vit_oracle import LayerNorm, QuickGELU

TEXT_GEOMETRY = {  # clip/model.py build_model: (embed_dim, context_length, vocab_size, width, heads, layers)
    "ViT-B/32": dict(embed_dim=512, context_length=77, vocab_size=49408, width=512, heads=8, layers=12),
    "ViT-B/16": dict(embed_dim=512, context_length=77, vocab_size=49408, width=512, heads=8, layers=12),
    "ViT-L/14": dict(embed_dim=768, context_length=77, vocab_size=49408, width=768, heads=12, layers=12),
}
SOT, EOT = 49406, 49407  # <|startoftext|>, <|endoftext|> of the 49 408-entry vocabulary (simple_tokenizer.py)


class MaskedBlock(nn.Module):
    """ResidualAttentionBlock with attn_mask (the text transformer passes the causal mask)."""

    def __init__(self, d, h, mask):
        super().__init__()
        self.attn = nn.MultiheadAttention(d, h)
        self.ln_1 = LayerNorm(d)
        self.mlp = nn.Sequential(OrderedDict([
            ("c_fc", nn.Linear(d, d * 4)), ("gelu", QuickGELU()), ("c_proj", nn.Linear(d * 4, d))]))
        self.ln_2 = LayerNorm(d)
        self.register_buffer("mask", mask, persistent=False)

    def forward(self, x):
        y = self.ln_1(x)
        x = x + self.attn(y, y, y, need_weights=False, attn_mask=self.mask[:x.shape[0], :x.shape[0]])[0]
        return x + self.mlp(self.ln_2(x))


class TextTransformerStack(nn.Module):
    def __init__(self, width, layers, heads, mask):
        super().__init__()
        self.resblocks = nn.Sequential(*[MaskedBlock(width, heads, mask) for _ in range(layers)])

    def forward(self, x):
        return self.resblocks(x)


class TextTower(nn.Module):
    """The text half of clip/model.py `CLIP` with the same attribute (= state_dict key) names."""

    def __init__(self, embed_dim, context_length, vocab_size, width, heads, layers):
        super().__init__()
        self.context_length = context_length
        mask = torch.empty(context_length, context_length).fill_(float("-inf")).triu_(1)  # build_attention_mask
        self.transformer = TextTransformerStack(width, layers, heads, mask)
        self.token_embedding = nn.Embedding(vocab_size, width)
        self.positional_embedding = nn.Parameter(torch.empty(context_length, width))
        self.ln_final = LayerNorm(width)
        self.text_projection = nn.Parameter(torch.empty(width, embed_dim))

    def forward(self, text):
        x = self.token_embedding(text) + self.positional_embedding     # [B, ctx, d]
        x = self.transformer(x.permute(1, 0, 2)).permute(1, 0, 2)
        x = self.ln_final(x)
        return x[torch.arange(x.shape[0]), text.argmax(dim=-1)] @ self.text_projection  # EOT row


def build_text(name="ViT-B/32", seed=5):
    """Seeded random init with openai's scales (initialize_parameters: token 0.02, positional 0.01, attn d^-0.5,
    out/c_proj d^-0.5 (2L)^-0.5, c_fc (2d)^-0.5, text_projection d^-0.5); LayerNorms and biases perturbed so that a
    wrong gamma/beta/bias wiring cannot pass."""
    g = TEXT_GEOMETRY[name]
    gen = torch.Generator().manual_seed(seed)
    m = TextTower(**g)
    d, L = g["width"], g["layers"]
    proj_std, attn_std, fc_std = (d ** -0.5) * ((2 * L) ** -0.5), d ** -0.5, (2 * d) ** -0.5
    with torch.no_grad():
        def rn(t, std):
            t.copy_(torch.randn(t.shape, generator=gen) * std)
        rn(m.token_embedding.weight, 0.02)
        rn(m.positional_embedding, 0.01)
        rn(m.text_projection, d ** -0.5)
        lns = [m.ln_final]
        for blk in m.transformer.resblocks:
            rn(blk.attn.in_proj_weight, attn_std)
            rn(blk.attn.in_proj_bias, 0.02)
            rn(blk.attn.out_proj.weight, proj_std)
            rn(blk.attn.out_proj.bias, 0.02)
            rn(blk.mlp.c_fc.weight, fc_std)
            rn(blk.mlp.c_fc.bias, 0.02)
            rn(blk.mlp.c_proj.weight, proj_std)
            rn(blk.mlp.c_proj.bias, 0.02)
            lns += [blk.ln_1, blk.ln_2]
        for ln in lns:
            ln.weight.copy_(1.0 + 0.1 * torch.randn(d, generator=gen))
            ln.bias.copy_(0.05 * torch.randn(d, generator=gen))
    return m.eval()


def synthetic_tokens(n, seed=0, context_length=77, vocab_size=49408):
    """Rows shaped like clip.tokenize output: SOT, 1..75 random word ids, EOT (the largest id), zero padding."""
    rng = np.random.default_rng(seed)
    out = np.zeros((n, context_length), dtype=np.int64)
    for i in range(n):
        k = int(rng.integers(1, context_length - 1)) if i else context_length - 2  # row 0 fills the whole context
        out[i, 0] = SOT
        out[i, 1:1 + k] = rng.integers(1, SOT, size=k)
        out[i, 1 + k] = EOT
    return torch.from_numpy(out)


@torch.no_grad()
def encode_text(tower, tokens, normalise=True):
    """Driven like embedding_service.py:169-180: encode_text -> numpy -> x / ||x||."""
    e = tower(tokens.long()).cpu().numpy()
    if normalise:
        e = e / np.linalg.norm(e, axis=-1, keepdims=True)
    return e
