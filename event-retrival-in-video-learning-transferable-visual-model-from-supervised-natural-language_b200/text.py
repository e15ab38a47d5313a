"""TextEncoder: the hand-written CLIP text tower behind `model.encode_text`.

Replaces openai/CLIP `CLIP.encode_text` (call site Backend/services/embedding_service.py:169-177: tokenize the
query, encode_text, L2-normalise).  Weights arrive as the openai-keyed state_dict of the whole CLIP model
(`token_embedding.weight`, `positional_embedding`, `transformer.resblocks.*`, `ln_final.*`, `text_projection`), are
repacked once to bf16 inside libevr.so, and a batch of token rows runs as the same sm_100a kernels as the vision
tower (tcgen05 GEMMs with fused epilogues, LayerNorm, attention core with the causal mask).
"""
import ctypes

import torch

from ._lib import HandleGuard, lib, check, ptr, current_stream_ptr, TextCfg


def text_weight_keys(layers):
    """State-dict keys in the order evr_text_create expects (include/evr.h)."""
    keys = ["token_embedding.weight", "positional_embedding"]
    for i in range(layers):
        b = f"transformer.resblocks.{i}."
        keys += [b + "ln_1.weight", b + "ln_1.bias", b + "attn.in_proj_weight", b + "attn.in_proj_bias",
                 b + "attn.out_proj.weight", b + "attn.out_proj.bias", b + "ln_2.weight", b + "ln_2.bias",
                 b + "mlp.c_fc.weight", b + "mlp.c_fc.bias", b + "mlp.c_proj.weight", b + "mlp.c_proj.bias"]
    return keys + ["ln_final.weight", "ln_final.bias", "text_projection"]


def text_geometry(sd):
    """clip/model.py build_model: geometry read off the checkpoint."""
    vocab, width = sd["token_embedding.weight"].shape
    layers = len({k.split(".")[2] for k in sd if k.startswith("transformer.resblocks.")})
    return dict(context_length=sd["positional_embedding"].shape[0], vocab_size=vocab, width=width, layers=layers,
                heads=width // 64, out_dim=sd["text_projection"].shape[1])


class TextEncoder:
    def __init__(self, cfg, weights, device, max_batch=64):
        self.cfg = dict(cfg)
        self.device = torch.device(device)
        self.max_batch = int(max_batch)
        self.out_dim = cfg["out_dim"]
        self.context_length = cfg["context_length"]
        self.vocab_size = cfg["vocab_size"]
        c = TextCfg(cfg["context_length"], cfg["vocab_size"], cfg["width"], cfg["layers"], cfg["heads"],
                    cfg["out_dim"])
        dev_w = [w.detach().to(device=self.device, dtype=torch.float32).contiguous() for w in weights]
        arr = (ctypes.c_void_p * len(dev_w))(*[w.data_ptr() for w in dev_w])
        handle = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            check(lib.evr_text_create(ctypes.byref(c), arr, len(dev_w), self.max_batch, ctypes.byref(handle)),
                  "evr_text_create")
            torch.cuda.synchronize()
        self._h = handle
        self._guard = HandleGuard()

    @classmethod
    def from_state_dict(cls, sd, device="cuda", max_batch=64):
        cfg = text_geometry(sd)
        keys = text_weight_keys(cfg["layers"])
        missing = [k for k in keys if k not in sd]
        if missing:
            raise KeyError(f"state_dict lacks {len(missing)} text-tower keys, e.g. {missing[:3]}")
        return cls(cfg, [sd[k] for k in keys], device, max_batch)

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            try:
                lib.evr_text_destroy(h)
            except Exception:
                pass
            self._h = None

    @torch.no_grad()
    def encode(self, tokens, normalise=False, out_dtype=torch.float32):
        """tokens: integer tensor [B, context_length] (what clip.tokenize emits) -> [B, out_dim] on the device."""
        if tokens.dim() != 2 or tokens.shape[1] != self.context_length:
            raise ValueError(f"tokens must be [B, {self.context_length}], got {tuple(tokens.shape)}")
        if tokens.dtype not in (torch.int32, torch.int64, torch.int16, torch.uint8):
            raise ValueError("tokens must be an integer tensor")
        if out_dtype not in (torch.float32, torch.float16):
            raise ValueError("out_dtype must be float32 or float16")
        t = tokens.to(self.device)
        B = t.shape[0]
        if B and (int(t.min()) < 0 or int(t.max()) >= self.vocab_size):
            raise ValueError(f"token ids must lie in [0, {self.vocab_size})")
        t = t.to(torch.int32).contiguous()
        out = torch.empty((B, self.out_dim), dtype=out_dtype, device=self.device)
        if B:
            with torch.cuda.device(self.device), self._guard:
                check(lib.evr_text_encode(self._h, ptr(t), B, ptr(out), 0 if out_dtype == torch.float32 else 1,
                                          1 if normalise else 0, current_stream_ptr()), "evr_text_encode")
        return out
