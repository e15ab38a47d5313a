"""CPU checker for the GPU preprocessing path (TEST INFRASTRUCTURE, never the shipped path).

The reference's preprocessing is clip._transform = torchvision Resize(n_px, BICUBIC) + CenterCrop(n_px) on PIL images
(/root/reference/Backend/services/embedding_service.py:475; the same constants at compare_models.py:387-391), so the
oracle here IS that code: PIL + torchvision are installed on both machines (`pil_transform_u8`).  PINNED: the
restatement below (`resample_reference_u8`, the integer arithmetic the CUDA kernels perform on the tables of
preprocess.plan) is checked bit-exactly against PIL in tests/test_preprocess.py.
"""
import numpy as np

PRECISION_BITS = 32 - 8 - 2  # libImaging/Resample.c


def pil_transform_u8(img, n_px=224):
    """uint8 [H, W, 3] -> uint8 [n_px, n_px, 3] through PIL itself (Resize bicubic + CenterCrop of clip._transform)."""
    from PIL import Image
    from torchvision import transforms as T
    pil = Image.fromarray(img)
    return np.asarray(T.CenterCrop(n_px)(T.Resize(n_px, interpolation=T.InterpolationMode.BICUBIC)(pil)))


def pil_stretch_u8(img, n_px=224):
    """transforms.Resize((n_px, n_px)) of the evaluation harness (compare_models.py:37, :387): bilinear, no crop."""
    from PIL import Image
    from torchvision import transforms as T
    return np.asarray(T.Resize((n_px, n_px))(Image.fromarray(img)))


def resample_reference_u8(img, p, n_px):
    """The two integer passes in numpy; `p` = preprocess.plan(w, h, n_px) (coefficient tables + row window)."""
    src = img.astype(np.int64)
    tmp = np.zeros((p["rows"], n_px, 3), dtype=np.uint8)
    half = 1 << (PRECISION_BITS - 1)
    for x in range(n_px):
        x0, n = p["hb"][x]
        acc = half + np.tensordot(src[p["y0"]:p["y0"] + p["rows"], x0:x0 + n, :], p["hk"][x, :n].astype(np.int64),
                                  axes=([1], [0]))
        tmp[:, x, :] = np.clip(acc >> PRECISION_BITS, 0, 255)
    out = np.zeros((n_px, n_px, 3), dtype=np.uint8)
    t = tmp.astype(np.int64)
    for y in range(n_px):
        r0, n = p["vb"][y]
        acc = half + np.tensordot(p["vk"][y, :n].astype(np.int64), t[r0 - p["y0"]:r0 - p["y0"] + n], axes=([0], [0]))
        out[y] = np.clip(acc >> PRECISION_BITS, 0, 255)
    return out
