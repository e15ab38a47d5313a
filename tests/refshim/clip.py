"""Test stand-in for the `clip` package (openai/CLIP), used ONLY to import the reference's service modules unmodified
in the authoring container (tests/golden/make_ref_service_golden.py, tests/test_query_surface.py).

The reference's `Backend/CLIP` submodule is absent, so its `import clip` cannot resolve.  This module
gives `load / tokenize / encode_text / encode_image` deterministic, weight-free behaviour: a query string maps to the
vector registered for it in TEXT_VECTORS (else to a vector seeded by the string's CRC), an image maps to a fixed random
projection of its 8x8 average-pooled pixels.  That is enough to run every line of embedding_service.py /
search_service.py / query_strategies.py that sits between the encoders and the ranked result -- the part of the path
whose arithmetic lives in the reference itself.  It says nothing about the towers (those are pinned separately).
"""
import zlib

import numpy as np
import torch

SOT, EOT, CONTEXT = 49406, 49407, 77
TEXT_VECTORS = {}   # query string -> np.float32 [D]; filled by the harness
DIM = 512


def available_models():
    return ["ViT-B/32"]


def tokenize(texts, context_length=CONTEXT, truncate=False):
    if isinstance(texts, str):
        texts = [texts]
    out = torch.zeros(len(texts), context_length, dtype=torch.long)
    for i, t in enumerate(texts):
        ids = [SOT] + [b + 1 for b in t.encode("utf-8")][:context_length - 2] + [EOT]
        out[i, :len(ids)] = torch.tensor(ids)
    return out


def _detokenize(row):
    ids = [int(x) for x in row if 0 < int(x) < 257]
    return bytes(i - 1 for i in ids).decode("utf-8", errors="replace")


class _Visual:
    output_dim = DIM
    input_resolution = 224


class FakeCLIP(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.visual = _Visual()
        self.logit_scale = torch.nn.Parameter(torch.tensor(2.6592))
        g = torch.Generator().manual_seed(4242)
        self.register_buffer("image_proj", torch.randn(192, DIM, generator=g))

    def encode_text(self, tokens):
        rows = []
        for row in tokens:
            text = _detokenize(row.tolist())
            vec = TEXT_VECTORS.get(text)
            if vec is None:
                vec = np.random.default_rng(zlib.crc32(text.encode("utf-8"))).standard_normal(DIM).astype(np.float32)
            rows.append(torch.from_numpy(np.asarray(vec, dtype=np.float32)))
        return torch.stack(rows)

    def encode_image(self, image):
        pooled = torch.nn.functional.adaptive_avg_pool2d(image.float(), 8).flatten(1)
        return pooled @ self.image_proj


def _preprocess(img):
    arr = np.asarray(img.convert("RGB").resize((224, 224)), dtype=np.float32) / 255.0
    return torch.from_numpy(arr).permute(2, 0, 1)


def load(name, device="cpu", jit=False, download_root=None):
    return FakeCLIP().to(device), _preprocess
