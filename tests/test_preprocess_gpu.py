"""CUDA resize + centre crop (evr_resize_crop_u8) against PIL itself: bit-exact uint8 output (integer arithmetic)."""
import importlib

import numpy as np
import pytest
import torch

from conftest import PKG
from oracle import preprocess_oracle as po
from oracle import vit_oracle as vo

pytestmark = pytest.mark.gpu


def _frames(b, h, w, seed):
    rng = np.random.default_rng(seed)
    img = rng.integers(0, 256, (b, h, w, 3), dtype=np.uint8)
    img[:, : h // 3] = np.where(rng.random((b, h // 3, w, 1)) < 0.5, 0, 255)  # hard edges exercise clip8
    return img


@pytest.mark.parametrize("h,w,b", [(360, 640, 5), (720, 1280, 2), (224, 224, 3), (300, 224, 2), (500, 333, 2),
                                   (100, 160, 4), (1080, 1920, 1), (37, 911, 2)])
def test_resize_crop_is_bit_exact(pkg, cuda, h, w, b):
    pre = pkg.preprocess.GpuPreprocess(224, device=cuda)
    img = _frames(b, h, w, h + w)
    out = pre(img)
    assert out.dtype == torch.uint8 and out.shape == (b, 224, 224, 3) and out.is_cuda
    out = out.cpu().numpy()
    for i in range(b):
        assert np.array_equal(out[i], po.pil_transform_u8(img[i], 224)), f"frame {i}"


def test_inputs_and_errors(pkg, cuda):
    from PIL import Image
    pre = pkg.preprocess.GpuPreprocess(224, device=cuda)
    img = _frames(1, 240, 320, 1)[0]
    ref = po.pil_transform_u8(img, 224)
    assert np.array_equal(pre(img)[0].cpu().numpy(), ref)                                   # single HWC frame
    assert np.array_equal(pre(Image.fromarray(img))[0].cpu().numpy(), ref)                  # PIL image
    assert np.array_equal(pre(torch.from_numpy(img).to(cuda))[0].cpu().numpy(), ref)        # device tensor
    assert pre(np.zeros((0, 240, 320, 3), np.uint8)).shape == (0, 224, 224, 3)
    with pytest.raises(ValueError):
        pre(np.zeros((1, 240, 320, 4), np.uint8))
    with pytest.raises(ValueError):
        pre(np.zeros((1, 240, 320, 3), np.float32))
    with pytest.raises(RuntimeError):
        pkg.preprocess.GpuPreprocess(224, device="cpu")
    other = pkg.preprocess.GpuPreprocess(336, device=cuda)(img)                             # ViT-L/14@336px geometry
    assert np.array_equal(other[0].cpu().numpy(), po.pil_transform_u8(img, 336))


def test_preprocess_feeds_the_encoder(pkg, cuda):
    """decoded frames -> GPU resize/crop -> encode (ToTensor + Normalize fused) == PIL preprocess -> fp32 oracle."""
    visual = vo.build_visual("ViT-B/32", seed=1)
    img = _frames(6, 360, 480, 77)
    pix = vo.normalise_u8(torch.from_numpy(np.stack([po.pil_transform_u8(f, 224) for f in img])))
    ref = vo.encode_image(visual, pix, normalise=True)
    enc = pkg.vit.VisionEncoder.from_state_dict(visual.state_dict(), "ViT-B/32", device=cuda, max_batch=8)
    out = enc.encode(pkg.preprocess.GpuPreprocess(224, device=cuda)(img), normalise=True).cpu().numpy()
    assert ((out * ref).sum(-1)).min() >= 0.999


@pytest.mark.parametrize("h,w", [(360, 640), (500, 333), (100, 160)])
def test_stretch_mode_is_bit_exact(pkg, cuda, h, w):
    """The evaluation harness' transform (compare_models.py:37: Resize((224, 224)), bilinear) on the device."""
    pre = pkg.preprocess.GpuPreprocess(224, device=cuda, mode="stretch")
    img = _frames(3, h, w, h * w)
    out = pre(img).cpu().numpy()
    for i in range(3):
        assert np.array_equal(out[i], po.pil_stretch_u8(img[i], 224))
