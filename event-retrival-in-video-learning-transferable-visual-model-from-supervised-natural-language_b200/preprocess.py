"""GPU image preprocessing: clip._transform's `Resize(n_px, BICUBIC)` + `CenterCrop(n_px)` on the device.

Replaces the per-frame PIL work of `self.preprocess(Image.open(p).convert('RGB'))`
(Backend/services/embedding_service.py:475, Backend/embedding.py:46).  The arithmetic is PIL's
8-bit `ImagingResample` (libImaging/Resample.c): separable convolution, horizontal pass then vertical pass, cubic
filter a = -0.5 with the support stretched by the down-scale factor (antialiasing), coefficients in 22-bit fixed point,
intermediate image in uint8.  The coefficient tables are built here on the host in float64 exactly as the C code does;
the two passes run as CUDA kernels in integers (csrc/preprocess.cu), so the result is bit-identical to PIL's
(tests/test_preprocess.py checks the tables + integer arithmetic against PIL itself on CPU,
tests/test_preprocess_gpu.py the kernels against PIL).  The output is uint8 [S, S, 3], which `VisionEncoder.encode` takes directly (ToTensor + Normalize fused).
"""
import ctypes
import functools

import numpy as np
import torch

from ._lib import lib, check, ptr, current_stream_ptr

PRECISION_BITS = 32 - 8 - 2  # Resample.c


def _bicubic(x):
    a = -0.5
    x = np.abs(x)
    return np.where(x < 1.0, ((a + 2.0) * x - (a + 3.0)) * x * x + 1.0,
                    np.where(x < 2.0, (((x - 5.0) * x + 8.0) * x - 4.0) * a, 0.0))


def _bilinear(x):
    x = np.abs(x)
    return np.where(x < 1.0, 1.0 - x, 0.0)


FILTERS = {"bicubic": (_bicubic, 2.0), "bilinear": (_bilinear, 1.0)}  # Resample.c BICUBIC / BILINEAR (filter, support)


@functools.lru_cache(maxsize=256)
def resample_coeffs(in_size, out_size, filt="bicubic"):
    """precompute_coeffs + normalize_coeffs_8bpc of Resample.c for one axis over the full input range [0, in_size).
    -> (bounds int32 [out, 2] = (xmin, count), kk int32 [out, ksize])."""
    kernel, support0 = FILTERS[filt]
    scale = filterscale = float(in_size) / out_size
    if filterscale < 1.0:
        filterscale = 1.0
    support = support0 * filterscale
    ksize = int(np.ceil(support)) * 2 + 1
    bounds = np.zeros((out_size, 2), dtype=np.int32)
    kk = np.zeros((out_size, ksize), dtype=np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        xmin = int(center - support + 0.5)
        if xmin < 0:
            xmin = 0
        xmax = int(center + support + 0.5)
        if xmax > in_size:
            xmax = in_size
        xmax -= xmin
        k = kernel((np.arange(xmax, dtype=np.float64) + xmin - center + 0.5) * ss)
        ww = 0.0
        for v in k:            # sequential double accumulation, as the C loop
            ww += v
        if ww != 0.0:
            k = k / ww
        fixed = np.where(k < 0, -0.5 + k * (1 << PRECISION_BITS), 0.5 + k * (1 << PRECISION_BITS))
        kk[xx, :xmax] = np.trunc(fixed).astype(np.int32)   # C (int) cast truncates toward zero
        bounds[xx] = (xmin, xmax)
    return bounds, kk


def resized_size(w, h, n_px):
    """torchvision Resize(int): the smaller edge becomes n_px, the other int(n_px * long / short)."""
    if w <= h:
        return n_px, int(n_px * h / w)
    return int(n_px * w / h), n_px


def crop_origin(w, h, n_px):
    """torchvision CenterCrop: (left, top) = round((size - n_px) / 2)."""
    return int(round((w - n_px) / 2.0)), int(round((h - n_px) / 2.0))


def plan(w, h, n_px, mode="clip"):
    """Everything the two device passes need for a w x h source: tables restricted to the output window.
    mode "clip":    Resize(n_px, BICUBIC) + CenterCrop(n_px)   (clip._transform; the service's ingest)
    mode "stretch": Resize((n_px, n_px)), torchvision's default BILINEAR, no crop
                    (the evaluation harness, content/Test_compare_model/compare_models.py:37, :387)"""
    if mode == "clip":
        rw, rh = resized_size(w, h, n_px)
        left, top = crop_origin(rw, rh, n_px)
        filt = "bicubic"
    elif mode == "stretch":
        rw = rh = n_px
        left = top = 0
        filt = "bilinear"
    else:
        raise ValueError(f"unknown preprocessing mode {mode!r}")
    hb, hk = resample_coeffs(w, rw, filt)
    vb, vk = resample_coeffs(h, rh, filt)
    hb, hk = hb[left:left + n_px], hk[left:left + n_px]
    vb, vk = vb[top:top + n_px], vk[top:top + n_px]
    y0 = int(vb[:, 0].min())
    y1 = int((vb[:, 0] + vb[:, 1]).max())
    return dict(rw=rw, rh=rh, left=left, top=top, hb=np.ascontiguousarray(hb), hk=np.ascontiguousarray(hk),
                vb=np.ascontiguousarray(vb), vk=np.ascontiguousarray(vk), y0=y0, rows=y1 - y0)


class GpuPreprocess:
    """Resize(n_px, BICUBIC) + CenterCrop(n_px) for batches of decoded frames of one size.

    __call__(frames): uint8 [B, H, W, 3] (numpy or torch, host or device; a single [H, W, 3] frame or a PIL image is
    accepted too) -> uint8 device tensor [B, n_px, n_px, 3], bit-identical to the PIL pipeline.  Feed it to
    `VisionEncoder.encode` / `model.encode_image`, which fuse ToTensor + Normalize."""

    def __init__(self, n_px=224, device="cuda", mode="clip"):
        self.n_px = int(n_px)
        self.mode = mode
        plan(n_px, n_px, n_px, mode)  # validates the mode
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("GpuPreprocess runs on a CUDA device only (the host path is clip.load's preprocess)")
        self._plans = {}

    def _plan(self, w, h):
        key = (w, h)
        if key not in self._plans:
            p = plan(w, h, self.n_px, self.mode)
            dev = {k: torch.from_numpy(p[k]).to(self.device) for k in ("hb", "hk", "vb", "vk")}
            self._plans[key] = (p, dev)
        return self._plans[key]

    @torch.no_grad()
    def __call__(self, frames):
        if hasattr(frames, "convert"):  # PIL image
            frames = np.array(frames.convert("RGB"))
        if isinstance(frames, np.ndarray):
            frames = torch.from_numpy(np.ascontiguousarray(frames))
        if frames.dim() == 3:
            frames = frames.unsqueeze(0)
        if frames.dtype != torch.uint8 or frames.dim() != 4 or frames.shape[-1] != 3:
            raise ValueError(f"frames must be uint8 [B, H, W, 3], got {frames.dtype} {tuple(frames.shape)}")
        src = frames.to(self.device).contiguous()
        B, H, W, _ = src.shape
        S = self.n_px
        out = torch.empty((B, S, S, 3), dtype=torch.uint8, device=self.device)
        if B == 0:
            return out
        p, dev = self._plan(W, H)
        done = 0
        with torch.cuda.device(self.device):
            while done < B:  # grid.z limit
                nb = min(B - done, 32768)
                tmp = torch.empty((nb, p["rows"], S, 3), dtype=torch.uint8, device=self.device)
                check(lib.evr_resize_crop_u8(ptr(src[done:done + nb]), nb, H, W, S, ptr(dev["hb"]), ptr(dev["hk"]),
                                             dev["hk"].shape[1], ptr(dev["vb"]), ptr(dev["vk"]), dev["vk"].shape[1],
                                             p["y0"], p["rows"], ptr(tmp), ptr(out[done:done + nb]),
                                             current_stream_ptr()), "evr_resize_crop_u8")
                done += nb
        return out


def jpeg_frame_size(blob):
    """(width, height) from the SOF0 / SOF1 header of a baseline JPEG byte string, None for anything else (progressive,
    lossless, arithmetic-coded, not a JPEG): those go through PIL.  A marker walk over the first segments only."""
    n = len(blob)
    if n < 4 or blob[0] != 0xFF or blob[1] != 0xD8:
        return None
    i = 2
    while i + 9 < n:
        if blob[i] != 0xFF:
            return None
        m = blob[i + 1]
        if m == 0xFF:                    # fill byte
            i += 1
            continue
        if m in (0xC0, 0xC1):            # baseline / extended sequential, Huffman
            return (blob[i + 7] << 8) | blob[i + 8], (blob[i + 5] << 8) | blob[i + 6]
        if 0xC2 <= m <= 0xCF and m not in (0xC4, 0xC8, 0xCC):
            return None                  # any other start-of-frame
        if m == 0xDA or m == 0xD9:       # scan data before a frame header
            return None
        if 0xD0 <= m <= 0xD7 or m == 0x01:
            i += 2
            continue
        i += 2 + ((blob[i + 2] << 8) | blob[i + 3])
    return None


class JpegDecoder:
    """JPEG files -> uint8 [B, H, W, 3] RGB frames in HBM through nvJPEG's batched decoder (`evr_jpeg_decode_batch`).

    Replaces `Image.open(p).convert('RGB')` of the ingest loop (Backend/services/embedding_service.py:472-475) when the
    service is created with `gpu_jpeg=True`.  The hardware JPEG engines are used when nvJPEG offers them for the stream
    (single-scan baseline JPEGs, what `cv2.imwrite` writes for key-frames, segment_video.py:26); otherwise nvJPEG's GPU
    Huffman path.  Pixels differ from libjpeg-turbo's by a few grey levels (IDCT / chroma up-sampling), so the `.npy`
    rows are not bit-identical to the PIL path's: embeddings agree to cosine >= 0.999 (tests/test_jpeg_gpu.py)."""

    def __init__(self, device="cuda", prefer_hardware=True):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("JpegDecoder runs on a CUDA device only (the host path is PIL)")
        self._h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            check(lib.evr_jpeg_create(1 if prefer_hardware else 0, ctypes.byref(self._h)), "evr_jpeg_create")
        self.backend = {3: "hardware", 2: "gpu_hybrid", 0: "default"}.get(int(lib.evr_jpeg_backend(self._h)), "other")

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            try:
                lib.evr_jpeg_destroy(h)
            except Exception:
                pass
            self._h = None

    def size(self, blob):
        """(width, height) of one JPEG byte string, from its header."""
        w, h, c = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        buf = (ctypes.c_ubyte * len(blob)).from_buffer_copy(blob)
        check(lib.evr_jpeg_info(self._h, buf, len(blob), ctypes.byref(w), ctypes.byref(h), ctypes.byref(c)),
              "evr_jpeg_info")
        return w.value, h.value

    @torch.no_grad()
    def decode(self, blobs, size=None):
        """blobs: list of JPEG byte strings of ONE image size -> uint8 device tensor [B, H, W, 3]."""
        n = len(blobs)
        if n == 0:
            return torch.empty((0, 0, 0, 3), dtype=torch.uint8, device=self.device)
        W, H = size if size is not None else self.size(blobs[0])
        keep = [b if isinstance(b, bytes) else bytes(b) for b in blobs]      # read in place, alive until the sync below
        ptrs = (ctypes.c_void_p * n)(*[ctypes.cast(ctypes.c_char_p(b), ctypes.c_void_p).value for b in keep])
        lens = (ctypes.c_size_t * n)(*[len(b) for b in blobs])
        out = torch.empty((n, H, W, 3), dtype=torch.uint8, device=self.device)
        with torch.cuda.device(self.device):
            check(lib.evr_jpeg_decode_batch(self._h, ptrs, lens, n, H, W, ptr(out), current_stream_ptr()),
                  "evr_jpeg_decode_batch")
            torch.cuda.current_stream().synchronize()   # nvJPEG reads the host bytes asynchronously: keep them alive
        return out

    def decode_files(self, paths):
        blobs = []
        for p in paths:
            with open(p, "rb") as f:
                blobs.append(f.read())
        return self.decode(blobs)
