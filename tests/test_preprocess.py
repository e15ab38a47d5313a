"""Preprocessing tables + integer arithmetic against PIL itself (CPU): clip._transform's Resize(224, BICUBIC) +
CenterCrop(224) must be reproduced bit for bit (the CUDA kernels perform exactly `resample_reference_u8`)."""
import importlib
import os

import numpy as np
import pytest

from conftest import PKG
from oracle import preprocess_oracle as po

SIZES = [(360, 640), (720, 1280), (224, 224), (300, 224), (500, 333), (100, 160), (1080, 1920), (225, 400), (37, 911)]


@pytest.mark.parametrize("h,w", SIZES)
def test_tables_and_integer_passes_equal_pil(h, w):
    pp = importlib.import_module(PKG + ".preprocess")
    rng = np.random.default_rng(h * 10007 + w)
    img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    img[: h // 3] = np.where(rng.random((h // 3, w, 1)) < 0.5, 0, 255)   # hard edges: the cubic lobes overshoot -> clip8
    ref = po.pil_transform_u8(img, 224)
    out = po.resample_reference_u8(img, pp.plan(w, h, 224), 224)
    assert np.array_equal(ref, out)


def test_geometry_rules():
    pp = importlib.import_module(PKG + ".preprocess")
    assert pp.resized_size(640, 360, 224) == (398, 224) and pp.resized_size(360, 640, 224) == (224, 398)
    assert pp.resized_size(333, 500, 224) == (224, 336)
    assert pp.crop_origin(398, 224, 224) == (87, 0) and pp.crop_origin(224, 337, 224) == (0, 56)  # round half to even
    b, k = pp.resample_coeffs(448, 224)
    assert k.shape == (224, 9) and (k.sum(1) - (1 << 22)).__abs__().max() <= 8   # rows sum to 1.0 in 22-bit fixed point
    assert (b[:, 0] >= 0).all() and (b[:, 0] + b[:, 1] <= 448).all()
    b, k = pp.resample_coeffs(224, 224)
    assert ((k != 0).sum(1) == 1).all() and (k.max(1) == 1 << 22).all()         # identity scale: one unit tap per row


@pytest.mark.parametrize("h,w", [(360, 640), (224, 224), (500, 333), (100, 160), (1080, 1920)])
def test_stretch_mode_equals_torchvision_resize(h, w):
    """compare_models.py's transform: Resize((224, 224)) = PIL bilinear with the support stretched by the scale."""
    pp = importlib.import_module(PKG + ".preprocess")
    rng = np.random.default_rng(h * 31 + w)
    img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    out = po.resample_reference_u8(img, pp.plan(w, h, 224, "stretch"), 224)
    assert np.array_equal(po.pil_stretch_u8(img, 224), out)
    with pytest.raises(ValueError):
        pp.plan(w, h, 224, "nearest")


def test_jpeg_frame_size_reads_sof_headers():
    """`jpeg_frame_size` (the run splitter of the nvJPEG ingest path) against PIL on the reference's own key-frames,
    plus the streams that must be refused (progressive, PNG, truncated)."""
    import io
    import importlib
    from PIL import Image
    from conftest import PKG
    size = importlib.import_module(PKG + ".preprocess").jpeg_frame_size
    d = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_frames")
    names = sorted(f for f in os.listdir(d) if f.endswith(".jpg"))
    assert names
    for n in names:
        blob = open(os.path.join(d, n), "rb").read()
        assert size(blob) == Image.open(io.BytesIO(blob)).size
    img = Image.fromarray((np.arange(96 * 160 * 3) % 251).astype(np.uint8).reshape(96, 160, 3))
    for kw, expect in ((dict(), (160, 96)), (dict(optimize=True, quality=40), (160, 96)),
                       (dict(progressive=True), None)):
        buf = io.BytesIO()
        img.save(buf, "JPEG", **kw)
        assert size(buf.getvalue()) == expect, kw
    buf = io.BytesIO()
    img.save(buf, "PNG")
    assert size(buf.getvalue()) is None
    assert size(b"") is None and size(b"\xff\xd8\xff") is None and size(blob[:12]) is None
