"""GPU parity of the hand-written CLIP vision tower (through the C ABI) against the fp32 CPU oracle
(oracle/vit_oracle.py, restatement of openai/CLIP clip/model.py) and of its tcgen05 GEMM building block against a
plain torch fp32 matmul.

Tolerance (north_star): bf16 operands with fp32 accumulation, fp32 residual stream and LayerNorm statistics =>
cosine(new, oracle fp32) >= 0.999 per embedding row; the GEMM itself must match an fp32 matmul of the same bf16
operands to fp32 round-off (bf16 output: one bf16 ulp)."""
import numpy as np
import pytest
import torch

from oracle import vit_oracle as vo

pytestmark = pytest.mark.gpu
MIN_COS = 0.999


def _gemm(pkg, A, W, bias, out, epi):
    lib, ptr = pkg._lib.lib, pkg._lib.ptr
    M, K = A.shape
    N = W.shape[0]
    pkg._lib.check(lib.evr_gemm_bf16(ptr(A), ptr(W), ptr(bias), ptr(out), M, N, K, epi, pkg._lib.current_stream_ptr()),
                   "evr_gemm_bf16")
    torch.cuda.synchronize()


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (1, 512, 768), (7, 512, 768), (200, 768, 768), (50, 2304, 768),
                                   (1600, 3072, 768), (12800, 768, 3072), (12544, 768, 3072), (514, 1024, 640),
                                   (12800, 2304, 768), (300, 32, 64), (129, 96, 128)])
@pytest.mark.parametrize("epi", [0, 1, 2, 3])
def test_gemm_epilogues(pkg, cuda, M, N, K, epi):
    g = torch.Generator().manual_seed(M * 31 + N * 7 + K + epi)
    A = (torch.randn(M, K, generator=g) * 0.5).to(cuda).bfloat16()
    W = (torch.randn(N, K, generator=g) * 0.05).to(cuda).bfloat16()
    bias = torch.randn(N, generator=g).to(cuda)
    ref = A.float() @ W.float().t() + bias
    if epi in (0, 1):
        out = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=cuda)
        if epi == 1:
            ref = ref * torch.sigmoid(1.702 * ref)  # QuickGELU, clip/model.py
        _gemm(pkg, A, W, bias, out, epi)
        tol = 2.0 ** -8 * ref.abs().clamp(min=1.0) + 1e-3  # one bf16 ulp
        assert torch.all((out.float() - ref).abs() <= tol)
    elif epi == 2:
        res = torch.randn(M, N, generator=g).to(cuda)
        out = res.clone()
        _gemm(pkg, A, W, bias, out, epi)
        assert torch.allclose(out, ref + res, atol=2e-4, rtol=1e-5)
    else:
        out = torch.full((M, N), float("nan"), dtype=torch.float32, device=cuda)
        _gemm(pkg, A, W, bias, out, epi)
        assert torch.allclose(out, ref, atol=2e-4, rtol=1e-5)


def test_gemm_without_bias(pkg, cuda):
    g = torch.Generator().manual_seed(3)
    A = torch.randn(64, 768, generator=g).to(cuda).bfloat16()
    W = (torch.randn(512, 768, generator=g) * 0.03).to(cuda).bfloat16()
    out = torch.empty((64, 512), dtype=torch.float32, device=cuda)
    _gemm(pkg, A, W, None, out, 3)
    assert torch.allclose(out, A.float() @ W.float().t(), atol=2e-4, rtol=1e-5)


def test_gemm_rejects_bad_shapes(pkg, cuda):
    A = torch.zeros((8, 100), dtype=torch.bfloat16, device=cuda)
    W = torch.zeros((64, 100), dtype=torch.bfloat16, device=cuda)
    out = torch.zeros((8, 64), dtype=torch.float32, device=cuda)
    with pytest.raises(pkg._lib.EvrError, match="K="):
        _gemm(pkg, A, W, None, out, 3)


def _cos(a, b):
    return (a * b).sum(-1) / np.linalg.norm(a, axis=-1) / np.linalg.norm(b, axis=-1)


@pytest.fixture(scope="module")
def b32():
    visual = vo.build_visual("ViT-B/32", seed=1)
    u8, pix = vo.synthetic_frames(37, seed=0)
    ref = vo.encode_image(visual, pix, normalise=False)
    return visual, u8, pix, ref


def test_vit_b32_matches_oracle(pkg, cuda, b32):
    visual, u8, pix, ref = b32
    enc = pkg.vit.VisionEncoder.from_state_dict(visual.state_dict(), "ViT-B/32", device=cuda, max_batch=16)
    out = enc.encode(pix.to(cuda)).cpu().numpy()  # 37 frames through max_batch 16: chunks of 16, 16, 5
    assert out.shape == (37, 512)
    cos = _cos(out, ref)
    assert cos.min() >= MIN_COS, cos.min()
    assert np.abs(out - ref).max() <= 0.02 * np.abs(ref).max()


def test_vit_batch_invariance(pkg, cuda, b32):
    visual, u8, pix, ref = b32
    enc = pkg.vit.VisionEncoder.from_state_dict(visual.state_dict(), "ViT-B/32", device=cuda, max_batch=64)
    whole = enc.encode(pix.to(cuda))
    one = torch.cat([enc.encode(pix[i:i + 1].to(cuda)) for i in (0, 5, 36)])
    assert torch.equal(whole[[0, 5, 36]], one)  # a frame's embedding does not depend on its batch


def test_vit_uint8_frames_and_normalised_fp16_output(pkg, cuda, b32):
    visual, u8, pix, ref = b32
    enc = pkg.vit.VisionEncoder.from_state_dict(visual.state_dict(), "ViT-B/32", device=cuda, max_batch=64)
    out = enc.encode(u8.to(cuda), normalise=True, out_dtype=torch.float16)  # ToTensor+Normalize fused on device
    assert out.dtype == torch.float16
    o = out.float().cpu().numpy()
    assert np.allclose(np.linalg.norm(o, axis=-1), 1.0, atol=2e-3)
    refn = ref / np.linalg.norm(ref, axis=-1, keepdims=True)  # embedding_service.py:502
    assert _cos(o, refn).min() >= MIN_COS


def test_vit_from_full_clip_state_dict_prefix(pkg, cuda, b32):
    visual, u8, pix, ref = b32
    sd = {"visual." + k: v for k, v in visual.state_dict().items()}
    sd["logit_scale"] = torch.tensor(4.6052)
    enc = pkg.vit.VisionEncoder.from_state_dict(sd, "ViT-B/32", device=cuda, max_batch=8, prefix="visual.")
    out = enc.encode(pix[:3].to(cuda)).cpu().numpy()
    assert _cos(out, ref[:3]).min() >= MIN_COS
    with pytest.raises(KeyError):
        pkg.vit.VisionEncoder.from_state_dict({}, "ViT-B/32", device=cuda)


@pytest.mark.parametrize("name,n", [("ViT-L/14", 3), ("ViT-B/16", 3)])
def test_other_geometries(pkg, cuda, name, n):
    visual = vo.build_visual(name, seed=2)
    u8, pix = vo.synthetic_frames(n, seed=4)
    ref = vo.encode_image(visual, pix, normalise=False)
    enc = pkg.vit.VisionEncoder.from_state_dict(visual.state_dict(), name, device=cuda, max_batch=2)
    out = enc.encode(pix.to(cuda)).cpu().numpy()
    assert out.shape == ref.shape
    assert _cos(out, ref).min() >= MIN_COS


def test_vit_argument_errors(pkg, cuda, b32):
    visual = b32[0]
    enc = pkg.vit.VisionEncoder.from_state_dict(visual.state_dict(), "ViT-B/32", device=cuda, max_batch=4)
    with pytest.raises(ValueError):
        enc.encode(torch.zeros((1, 3, 200, 200), device=cuda))
    with pytest.raises(ValueError):
        enc.encode(torch.zeros((1, 224, 224, 4), dtype=torch.uint8, device=cuda))
    assert enc.encode(torch.zeros((0, 3, 224, 224), device=cuda)).shape == (0, 512)


@pytest.mark.parametrize("name,n_img", [("ViT-B/32", 160), ("ViT-L/14", 12)])
def test_retrieval_metrics_unchanged(pkg, cuda, name, n_img):
    """BASELINE config 5 in small: encode -> index -> query, R@K / MRR by compare_models.py:1004-1090's definitions.
    The metrics of the GPU embeddings must equal those of the fp32 oracle embeddings on the same synthetic frames and
    captions (5 noisy copies of each oracle image embedding); a rank may move only where two similarities are closer
    than the bf16 tower's error (random-init towers give embeddings with mean pairwise cosine 0.98, so gaps are small):
    R@K may differ by max(2 captions, 0.01) and MRR by 0.01 (stated tolerance).  The t2i
    ranks are also recomputed through the fused scan (FrameStore.search over the GPU-encoded store)."""
    from oracle import search_oracle as so
    visual = vo.build_visual(name, seed=3)
    u8, pix = vo.synthetic_frames(n_img, seed=6)
    ref = vo.encode_image(visual, pix, normalise=True)
    enc = pkg.vit.VisionEncoder.from_state_dict(visual.state_dict(), name, device=cuda, max_batch=64)
    emb_dev = enc.encode(pix.to(cuda), normalise=True)
    emb = emb_dev.cpu().numpy()
    assert _cos(emb, ref).min() >= MIN_COS
    rng = np.random.default_rng(8)
    caps = 5
    # random-init towers give highly correlated image embeddings: the noise is scaled to the spread between them
    spread = np.linalg.norm(ref - ref.mean(0), axis=-1).mean()
    txt = np.repeat(ref, caps, axis=0) + rng.standard_normal((n_img * caps, ref.shape[1])).astype(np.float32) \
        * (7.0 * spread / np.sqrt(ref.shape[1]))
    txt /= np.linalg.norm(txt, axis=-1, keepdims=True)
    m_ref, m_gpu = so.t2i_i2t_metrics(ref, txt, caps), so.t2i_i2t_metrics(emb, txt, caps)
    assert 0.05 < m_ref["t2i"]["R@1"] < 1.0 or n_img < 20      # the task is not degenerate
    for side, n in (("t2i", n_img * caps), ("i2t", n_img)):
        for k in ("R@1", "R@5", "R@10"):
            assert abs(m_ref[side][k] - m_gpu[side][k]) <= max(2.0 / n, 0.01) + 1e-9, (side, k, m_ref[side], m_gpu[side])
        assert abs(m_ref[side]["MRR"] - m_gpu[side]["MRR"]) <= 0.01
    # t2i through the product path: captions are the queries, the GPU-encoded frames are the store
    store = pkg.search.FrameStore.from_tensor(emb_dev, device=cuda)
    k = min(10, n_img)
    sc, idx = store.search(torch.from_numpy(txt).to(cuda), k)
    idx = idx.cpu().numpy()
    gt = np.arange(n_img * caps) // caps
    for kk in (1, 5, 10):
        if kk <= k:
            r_at_k = float(np.mean((idx[:, :kk] == gt[:, None]).any(1)))
            assert abs(r_at_k - m_ref["t2i"][f"R@{kk}"]) <= max(2.0 / (n_img * caps), 0.01) + 1e-9, (kk, r_at_k, m_ref["t2i"])


def test_small_batch_graph_replay_is_identical(pkg, cuda, b32):
    """Batches <= 32 run as a captured CUDA graph from their third call on (plain, capture, replay...): same bits as
    the plain pass, fresh inputs are honoured, and the launch counter advances by the same number of kernels."""
    visual, u8, pix, ref = b32
    enc = pkg.vit.VisionEncoder.from_state_dict(visual.state_dict(), "ViT-B/32", device=cuda, max_batch=64)
    a, b = pix[:5].to(cuda), pix[5:10].to(cuda)
    lib = pkg._lib.lib
    c0 = lib.evr_launch_count()
    plain = enc.encode(a).clone()
    c1 = lib.evr_launch_count()
    captured = enc.encode(a).clone()       # capture + first replay
    c2 = lib.evr_launch_count()
    replay_b = enc.encode(b).clone()       # replay on other pixels
    replay_a = enc.encode(a).clone()
    c3 = lib.evr_launch_count()
    assert torch.equal(plain, captured) and torch.equal(plain, replay_a)
    assert not torch.equal(replay_a, replay_b)
    assert _cos(replay_b.cpu().numpy(), ref[5:10]).min() >= MIN_COS
    assert c1 - c0 == c2 - c1 and c3 - c2 == 2 * (c1 - c0)
    # a side stream works too (the graph is launched on the caller's stream)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        on_side = enc.encode(b).clone()
    s.synchronize()
    assert torch.equal(on_side, replay_b)


def test_profile_modes_do_not_change_results(pkg, cuda, b32):
    """evr_profile_enable(1) brackets launches with events, (2) lets the GEMM / LayerNorm / attention kernels stamp
    %globaltimer: same outputs, the same number of launches reported by both, positive times, and a timeline whose
    kernels do not overlap (every stamp is taken after griddepcontrol.wait)."""
    visual, u8, pix, ref = b32
    enc = pkg.vit.VisionEncoder.from_state_dict(visual.state_dict(), "ViT-B/32", device=cuda, max_batch=64)
    x = pix[:40].to(cuda)  # above the captured-graph batch limit: plain launches in every mode
    base = enc.encode(x).clone()
    counts = {}
    for mode in (1, 2):
        pkg._lib.profile_enable(mode)
        out = enc.encode(x).clone()
        torch.cuda.synchronize()
        fam = pkg._lib.profile_read()
        timeline = pkg._lib.profile_timeline() if mode == 2 else None
        pkg._lib.profile_enable(False)
        assert torch.equal(out, base)
        ms, n = fam["gemm_tc"]
        assert n == 2 + 4 * 12 and ms > 0.0, (mode, fam)  # patch + projection + 4 per block
        counts[mode] = fam
    for f, n in (("layernorm", 2 + 2 * 12), ("attention", 12)):
        assert counts[1][f][1] == n and counts[2][f][1] == n and counts[2][f][0] > 0.0, (f, counts)
    assert counts[2]["vit_misc"][1] == 0                     # the small kernels carry no device stamp
    assert len(timeline) == 2 + 4 * 12 + 2 + 2 * 12 + 12
    ordered = sorted(timeline, key=lambda t: t[1])
    assert all(a[2] <= b[1] for a, b in zip(ordered, ordered[1:]))   # one kernel of the chain at a time


def test_gemm_bits_do_not_depend_on_tile_width(pkg, cuda, monkeypatch):
    """The tile chooser (width 128 / 192 / 256, tail strips) is free to change with the batch: an output element is
    the same K-ordered sum in every tiling, so the bits must not move (this is what keeps a frame's embedding
    independent of the batch it was encoded in)."""
    g = torch.Generator().manual_seed(11)
    M, N, K = 3000, 768, 768
    A = (torch.randn(M, K, generator=g) * 0.5).to(cuda).bfloat16()
    W = (torch.randn(N, K, generator=g) * 0.05).to(cuda).bfloat16()
    bias = torch.randn(N, generator=g).to(cuda)
    outs = []
    for bn in ("128", "192", "256"):
        monkeypatch.setenv("EVR_GEMM_BN", bn)
        out = torch.empty((M, N), dtype=torch.float32, device=cuda)
        _gemm(pkg, A, W, bias, out, 3)
        outs.append(out)
    monkeypatch.delenv("EVR_GEMM_BN")
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
