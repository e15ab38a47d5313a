"""Developer probe: runs each kernel family on cuda:0 and prints error magnitudes (no asserts).
    python scripts/gpu_probe.py [gemm] [ln] [attn] [vit] [search]
"""
import os
import sys
import time
import traceback

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
from oracle import search_oracle, vit_oracle  # noqa: E402
lib, check, ptr, sp = pkg._lib.lib, pkg._lib.check, pkg._lib.ptr, pkg._lib.current_stream_ptr
dev = torch.device("cuda:0")
which = set(sys.argv[1:]) or {"gemm", "vit", "search"}


def gemm_case(M, N, K, epi):
    g = torch.Generator(device="cpu").manual_seed(M + N + K + epi)
    A = (torch.randn(M, K, generator=g) * 0.5).to(dev).bfloat16()
    W = (torch.randn(N, K, generator=g) * 0.05).to(dev).bfloat16()
    bias = torch.randn(N, generator=g).to(dev)
    ref = A.float() @ W.float().t() + bias
    if epi == 0:
        out = torch.zeros(M, N, dtype=torch.bfloat16, device=dev)
    elif epi == 1:
        out = torch.zeros(M, N, dtype=torch.bfloat16, device=dev)
        ref = ref * torch.sigmoid(1.702 * ref)
    elif epi == 2:
        res = torch.randn(M, N, generator=g).to(dev)
        out = res.clone()
        ref = ref + res
    else:
        out = torch.zeros(M, N, dtype=torch.float32, device=dev)
    check(lib.evr_gemm_bf16(ptr(A), ptr(W), ptr(bias), ptr(out), M, N, K, epi, sp()), "gemm")
    torch.cuda.synchronize()
    err = (out.float() - ref).abs().max().item()
    rel = err / ref.abs().max().item()
    print(f"gemm M={M} N={N} K={K} epi={epi}: max abs err {err:.4e} rel {rel:.3e}", flush=True)


def run(name, fn):
    try:
        fn()
    except Exception:
        print(f"[{name}] FAILED", flush=True)
        traceback.print_exc()
        try:
            torch.cuda.synchronize()
        except Exception as e:
            print("device state:", e, flush=True)
            sys.exit(3)


if "gemm" in which:
    def _g():
        gemm_case(128, 128, 64, 3)
        gemm_case(128, 128, 256, 3)
        gemm_case(256, 256, 768, 3)
        gemm_case(200, 768, 768, 0)
        gemm_case(12800, 2304, 768, 0)
        gemm_case(12800, 3072, 768, 1)
        gemm_case(12800, 768, 3072, 2)
        gemm_case(7, 512, 768, 3)
    run("gemm", _g)

if "search" in which:
    def _s():
        for (N, Q, k) in [(4096, 8, 10), (1000, 3, 100), (50000, 130, 100), (300000, 1, 100), (20000, 5, 300)]:
            rng = np.random.default_rng(N)
            raw = rng.standard_normal((N, 512)).astype(np.float32)
            q = rng.standard_normal((Q, 512)).astype(np.float32)
            q /= np.linalg.norm(q, axis=-1, keepdims=True)
            store = pkg.search.FrameStore.from_numpy(raw, device=dev)
            emb = search_oracle.normalise_store(raw)
            err_norm = (store.rows.float().cpu().numpy() - emb).__abs__().max()
            t0 = time.time()
            sc, idx = store.search(torch.from_numpy(q).to(dev), k)
            torch.cuda.synchronize()
            dt = time.time() - t0
            sc, idx = sc.cpu().numpy(), idx.cpu().numpy()
            ref_i, ref_s = search_oracle.batched_topk(emb, q, k)
            ok = sum(search_oracle.topk_equal_up_to_ties(idx[j], sc[j], ref_i[j], ref_s[j], 2.0 ** -10) for j in range(Q))
            exact = float((idx == ref_i).mean())
            print(f"search N={N} Q={Q} k={k}: store err {err_norm:.2e}, parity {ok}/{Q}, exact idx frac {exact:.4f}, "
                  f"max score err {np.abs(sc - ref_s).max():.2e}, {dt*1e3:.1f} ms", flush=True)
    run("search", _s)

if "vit" in which:
    def _v():
        visual = vit_oracle.build_visual("ViT-B/32", seed=1)
        u8, pix = vit_oracle.synthetic_frames(6, seed=0)
        ref = vit_oracle.encode_image(visual, pix, normalise=False)
        enc = pkg.vit.VisionEncoder.from_state_dict(visual.state_dict(), "ViT-B/32", device=dev, max_batch=4)
        out = enc.encode(pix.to(dev), normalise=False).cpu().numpy()
        cos = (out * ref).sum(-1) / np.linalg.norm(out, axis=-1) / np.linalg.norm(ref, axis=-1)
        print("vit fp32-pixels cosine:", cos, "max abs", np.abs(out - ref).max(), "ref scale", np.abs(ref).max(), flush=True)
        out8 = enc.encode(u8.to(dev), normalise=True).cpu().numpy()
        refn = ref / np.linalg.norm(ref, axis=-1, keepdims=True)
        print("vit uint8 cosine:", (out8 * refn).sum(-1), flush=True)
    run("vit", _v)
print("probe done", flush=True)
