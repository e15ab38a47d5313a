"""VisionEncoder: the hand-written CLIP vision tower behind `model.encode_image`.

Replaces openai/CLIP `CLIP.encode_image -> VisionTransformer.forward` (call sites Backend/embedding.py:49,
Backend/services/embedding_service.py:41,414,490).  Weights arrive as an openai-keyed state_dict, are repacked
once to bf16 inside libevr.so, and every batch runs as a fixed sequence of sm_100a kernels (tcgen05 GEMMs with
fused bias / QuickGELU / residual epilogues, LayerNorm, attention core).
"""
import ctypes

import torch

from ._lib import HandleGuard, lib, check, ptr, current_stream_ptr, VitCfg

GEOMETRY = {  # clip/model.py build_model
    "ViT-B/32": dict(image_size=224, patch=32, width=768, layers=12, heads=12, out_dim=512),
    "ViT-B/16": dict(image_size=224, patch=16, width=768, layers=12, heads=12, out_dim=512),
    "ViT-L/14": dict(image_size=224, patch=14, width=1024, layers=24, heads=16, out_dim=768),
}


def weight_keys(layers, prefix=""):
    """openai/CLIP VisionTransformer state_dict keys in the order evr_vit_create expects (include/evr.h)."""
    keys = ["conv1.weight", "class_embedding", "positional_embedding", "ln_pre.weight", "ln_pre.bias"]
    for i in range(layers):
        b = f"transformer.resblocks.{i}."
        keys += [b + "ln_1.weight", b + "ln_1.bias", b + "attn.in_proj_weight", b + "attn.in_proj_bias",
                 b + "attn.out_proj.weight", b + "attn.out_proj.bias", b + "ln_2.weight", b + "ln_2.bias",
                 b + "mlp.c_fc.weight", b + "mlp.c_fc.bias", b + "mlp.c_proj.weight", b + "mlp.c_proj.bias"]
    keys += ["ln_post.weight", "ln_post.bias", "proj"]
    return [prefix + k for k in keys]


class VisionEncoder:
    def __init__(self, cfg, weights, device, max_batch=256):
        self.cfg = dict(cfg)
        self.device = torch.device(device)
        self.max_batch = int(max_batch)
        self.out_dim = cfg["out_dim"]
        self.image_size = cfg["image_size"]
        c = VitCfg(cfg["image_size"], cfg["patch"], cfg["width"], cfg["layers"], cfg["heads"], cfg["out_dim"])
        dev_w = [w.detach().to(device=self.device, dtype=torch.float32).contiguous() for w in weights]
        arr = (ctypes.c_void_p * len(dev_w))(*[w.data_ptr() for w in dev_w])
        handle = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            check(lib.evr_vit_create(ctypes.byref(c), arr, len(dev_w), self.max_batch, ctypes.byref(handle)),
                  "evr_vit_create")
            torch.cuda.synchronize()
        self._h = handle
        self._guard = HandleGuard()

    @classmethod
    def from_state_dict(cls, sd, name="ViT-B/32", device="cuda", max_batch=256, prefix=""):
        """sd: openai-keyed state_dict of the visual tower (prefix "" ) or of the whole CLIP (prefix "visual.")."""
        cfg = GEOMETRY[name]
        keys = weight_keys(cfg["layers"], prefix)
        missing = [k for k in keys if k not in sd]
        if missing:
            raise KeyError(f"state_dict lacks {len(missing)} vision-tower keys, e.g. {missing[:3]}")
        return cls(cfg, [sd[k] for k in keys], device, max_batch)

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            try:
                lib.evr_vit_destroy(h)
            except Exception:
                pass
            self._h = None

    @torch.no_grad()
    def encode(self, pixels, normalise=False, out_dtype=torch.float32):
        """pixels: [B,3,S,S] float (already normalised, as clip's preprocess emits) or [B,S,S,3] uint8 RGB.
        Returns [B, out_dim] on the device (fp32 or fp16)."""
        S = self.image_size
        if pixels.dtype == torch.uint8:
            if pixels.dim() != 4 or tuple(pixels.shape[1:]) != (S, S, 3):
                raise ValueError(f"uint8 frames must be [B,{S},{S},3], got {tuple(pixels.shape)}")
            pix, pix_dtype = pixels.to(self.device).contiguous(), 1
        else:
            if pixels.dim() != 4 or tuple(pixels.shape[1:]) != (3, S, S):
                raise ValueError(f"float frames must be [B,3,{S},{S}], got {tuple(pixels.shape)}")
            pix, pix_dtype = pixels.to(device=self.device, dtype=torch.float32).contiguous(), 0
        if out_dtype not in (torch.float32, torch.float16):
            raise ValueError("out_dtype must be float32 or float16")
        B = pix.shape[0]
        out = torch.empty((B, self.out_dim), dtype=out_dtype, device=self.device)
        if B:
            with torch.cuda.device(self.device), self._guard:
                check(lib.evr_vit_encode(self._h, ptr(pix), pix_dtype, B, ptr(out),
                                         0 if out_dtype == torch.float32 else 1, 1 if normalise else 0,
                                         current_stream_ptr()), "evr_vit_encode")
        return out

    @torch.no_grad()
    def encode_stream(self, host_batches, normalise=True, out_dtype=torch.float32):
        """Pipelined host->host encode: for every host batch ([B,S,S,3] uint8 or [B,3,S,S] float, ideally pinned)
        yields a host tensor [B, out_dim].  The H2D copy of batch i+1 runs on a side stream while batch i is in the
        tower, and batch i is handed back once its embeddings have landed in pinned memory -- the device-side shape
        of the loop at Backend/services/embedding_service.py:467-493 (stack -> .to(device) -> encode_image ->
        .cpu().numpy()), without its per-batch stall."""
        dev = self.device
        with torch.cuda.device(dev):
            comp = torch.cuda.current_stream(dev)
            copy = self._copy_stream = getattr(self, "_copy_stream", None) or torch.cuda.Stream(device=dev)
            stage = [None, None]
            out_host = [None, None]
            ev_in = [torch.cuda.Event(), torch.cuda.Event()]
            ev_free = [None, None]
            ev_out = [torch.cuda.Event(), torch.cuda.Event()]
            pending = None
            for i, hb in enumerate(host_batches):
                slot = i & 1
                if isinstance(hb, torch.Tensor) and hb.is_cuda:
                    raise ValueError("encode_stream takes host batches; use encode() for device tensors")
                hb = torch.as_tensor(hb)
                B = hb.shape[0]
                if hb.dtype != torch.uint8:
                    hb = hb.to(torch.float32)
                if stage[slot] is None or stage[slot].shape[0] < B or stage[slot].dtype != hb.dtype \
                        or stage[slot].shape[1:] != hb.shape[1:]:
                    stage[slot] = torch.empty((max(B, 1),) + tuple(hb.shape[1:]), dtype=hb.dtype, device=dev)
                    ev_free[slot] = None
                    # the caching allocator may hand back a block that compute-stream work still in flight is
                    # reading (a previous encode's temporary, an embedding whose D2H copy is pending): the copy
                    # stream must not write it before that work has drained
                    copy.wait_stream(comp)
                if out_host[slot] is None or out_host[slot].shape[0] < B or out_host[slot].dtype != out_dtype:
                    out_host[slot] = torch.empty((max(B, 1), self.out_dim), dtype=out_dtype).pin_memory()
                with torch.cuda.stream(copy):
                    if ev_free[slot] is not None:
                        copy.wait_event(ev_free[slot])  # the tower has finished reading this staging slot
                    stage[slot][:B].copy_(hb, non_blocking=True)
                    ev_in[slot].record(copy)
                comp.wait_event(ev_in[slot])
                emb = self.encode(stage[slot][:B], normalise=normalise, out_dtype=out_dtype)
                ev_free[slot] = torch.cuda.Event()
                ev_free[slot].record(comp)
                if pending is not None:  # slot^1's embeddings: wait, hand back, then reuse its pinned buffer
                    pslot, pB = pending
                    ev_out[pslot].synchronize()
                    yield out_host[pslot][:pB].clone()
                out_host[slot][:B].copy_(emb, non_blocking=True)
                ev_out[slot].record(comp)
                pending = (slot, B)
            if pending is not None:
                pslot, pB = pending
                ev_out[pslot].synchronize()
                yield out_host[pslot][:pB].clone()
