"""LayerNorm of the residual stream (building block behind `evr_layernorm`) against PyTorch's fp32 LayerNorm.

openai/CLIP clip/model.py `LayerNorm`: statistics in fp32, eps 1e-5.  The reference here is torch's LayerNorm in
float64.  Tolerance: fp32 output within 2e-5 * (1 + |y|) (fp32 sums over up to 1024 values with an outlier channel of
40 sigma); bf16 output within one bf16 ulp (2^-7 relative) of the exact value.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(pkg, x, g, b, rows, d, out_bf16, in_stride=None, out=None, out_stride=None):
    lib, check, ptr, stream = pkg._lib.lib, pkg._lib.check, pkg._lib.ptr, pkg._lib.current_stream_ptr
    if out is None:
        out = torch.full((rows, d), float("nan"), dtype=torch.bfloat16 if out_bf16 else torch.float32, device=x.device)
    check(lib.evr_layernorm(ptr(x), in_stride or d, ptr(out), out_stride or d, int(out_bf16), ptr(g), ptr(b), rows, d,
                            stream()), "evr_layernorm")
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("rows,d", [(1, 768), (50, 768), (9472, 768), (9480, 768), (12800, 768), (12801, 512),
                                    (32896, 1024), (19712, 512)])
def test_layernorm_matches_torch(pkg, rows, d):
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(rows + d)
    x = torch.randn((rows, d), device=dev, generator=g) * 3.0 + torch.randn((rows, 1), device=dev, generator=g) * 5.0
    x[:, 7] += 40.0                                    # an outlier channel, as CLIP's residual stream has
    gamma = torch.rand((d,), device=dev, generator=g) + 0.5
    beta = torch.randn((d,), device=dev, generator=g)
    ref = torch.nn.functional.layer_norm(x.double(), (d,), gamma.double(), beta.double(), 1e-5).float()
    y32 = _run(pkg, x, gamma, beta, rows, d, False)
    assert torch.all((y32 - ref).abs() <= 2e-5 * (1 + ref.abs())), float((y32 - ref).abs().max())
    y16 = _run(pkg, x, gamma, beta, rows, d, True)
    ulp = ref.abs() * 2.0 ** -7 + 2e-5                     # one bf16 ulp is at most 2^-7 relative
    assert torch.all((y16.float() - ref).abs() <= ulp), float(((y16.float() - ref).abs() / ulp).max())


def test_layernorm_in_place_and_strided(pkg):
    """ln_pre runs in place on the fp32 stream; ln_post reads one row per image (stride L * d)."""
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(3)
    rows, d, L = 12800, 768, 50
    x = torch.randn((rows, d), device=dev, generator=g)
    gamma, beta = torch.rand((d,), device=dev, generator=g) + 0.5, torch.randn((d,), device=dev, generator=g)
    ref = torch.nn.functional.layer_norm(x, (d,), gamma, beta, 1e-5)
    y = x.clone()
    _run(pkg, y, gamma, beta, rows, d, False, out=y)
    assert torch.all((y - ref).abs() <= 2e-5 * (1 + ref.abs()))
    cls = _run(pkg, x, gamma, beta, rows // L, d, True, in_stride=L * d)
    assert torch.all((cls.float() - ref[::L]).abs() <= ref[::L].abs() * 2.0 ** -7 + 2e-5)


def test_layernorm_rejects_unsupported_width(pkg):
    dev = torch.device("cuda:0")
    x = torch.zeros((4, 640), device=dev)
    with pytest.raises(pkg._lib.EvrError):
        _run(pkg, x, x[0], x[0], 4, 640, False)
