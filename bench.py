#!/usr/bin/env python
"""bench.py -- headline benchmark of the retrieval hot path on B200 (contract: see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload both|encode|search]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

Primary line (BASELINE.json configs[1]): ViT-B/32 key-frame encode, batches of 256 synthetic frames per GPU, bf16 tensor
cores.  A step = one pass over `--inner` (64) consecutive batches of 256 frames, as a folder ingest walks its frame
list batch by batch (embedding_service.py:467); K = 20 steps are seconds of sustained work, so the denominators are the
SUSTAINED peaks.  value = frames/s over all ranks with the uint8 frames already resident in HBM; e2e = the same through
the public host API (pinned uint8 frames -> H2D -> tower -> D2H of the embeddings; at N > 1 through the data-parallel
ingest: rank slices of one frame list into one shared .npy) inside the timed region.
Nested "search" object (BASELINE.json configs[2]; configs[3] for N > 1): fused cosine scan + top-100,
10 000 queries x 10 M x 512 fp16 rows on one GPU (100 M rows row-sharded + NCCL all-gather + merge for N > 1), plus
the HBM-bound small-query-batch regime; at N = 1 the 100 M-row store is also run on the single GPU ("search_100m":
the N = 1 point of the strong-scaling curve).
Every line carries a `parity` object: checks of the timed configuration's own results, outside the timed regions.

--impl reference times the reference's own CPU path (its algorithm restated in oracle/, see DESIGN.md: the `clip`
dependency of the reference is an absent submodule) on the host cores, on a bounded sample of the same workload.
"""
import argparse
import contextlib
import importlib.util
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "frames/s ViT-B/32 encode"
SEARCH_METRIC = "queries/s top-100 over 10M-frame store"
VIT_NAME = "ViT-B/32"


# ----------------------------------------------------------------------------------------------- helpers
def ncu_traffic(kernel):
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture (profiles/traffic.json:
    dram__bytes_read.sum + dram__bytes_write.sum averaged over the captured launches); None when there is no capture."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f)[kernel]["dram_bytes_per_launch"]
    except (OSError, KeyError, ValueError):
        return None


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "tf_burst": d["bf16_tflops"], "tf_sustained": d["bf16_tflops_sustained"],
                "source": "measured (MEASURED_PEAKS.json)"}
    return {"hbm_gbs": 6650.0, "tf_burst": 1590.0, "tf_sustained": 1400.0, "source": "fallback (B200_PROFILING.md)"}


def vit_flops_per_frame(name):
    """(total, gemm-family) algorithmic FLOPs per frame; SURVEY.md 8d / BASELINE.md section 3."""
    g = {"ViT-B/32": (224, 32, 768, 12, 12, 512), "ViT-L/14": (224, 14, 1024, 24, 16, 768)}[name]
    S, p, d, layers, heads, out = g
    P = (S // p) ** 2
    L = P + 1
    patch = 2.0 * P * 3 * p * p * d
    per_layer_gemm = 2.0 * L * d * (3 * d + d + 4 * d + 4 * d)
    attn = 2.0 * 2.0 * heads * L * L * (d // heads)
    proj = 2.0 * d * out
    gemm = patch + layers * per_layer_gemm + proj
    return gemm + layers * attn, gemm


def eager_module():
    """scripts/torch_eager_baseline.py: the stock-PyTorch tower (library baseline; in fp32 also the parity checker)."""
    spec = importlib.util.spec_from_file_location("torch_eager_baseline",
                                                  os.path.join(ROOT, "scripts", "torch_eager_baseline.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled every 200 ms during the timed region (B200_PROFILING.md)."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "200"], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, power, reasons, mx = [], [], set(), None
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx = float(parts[1])
                power.append(float(parts[2]))
            except ValueError:
                continue
            for nm, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.path)
        if sm:
            # samples under load = those drawing at least 60 % of the highest power seen (the first and last
            # samples fall outside the timed region)
            cut = 0.6 * max(power) if power else 0
            busy = [c for c, w in zip(sm, power) if w >= cut] or sm
            out.update(sm_mhz=statistics.median(busy), sm_min_mhz=min(busy), sm_max_mhz=mx, reasons=sorted(reasons),
                       samples=len(sm), samples_under_load=len(busy), power_w_max=max(power) if power else None)
        return out


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def reduce_over_ranks(x, device, world, op="max"):
    import torch
    import torch.distributed as dist
    if world == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op={"max": dist.ReduceOp.MAX, "min": dist.ReduceOp.MIN, "sum": dist.ReduceOp.SUM}[op])
    return float(t.item())


def max_over_ranks(x, device, world):
    return reduce_over_ranks(x, device, world, "max")


def barrier_sync(world):
    import torch
    import torch.distributed as dist
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def chunked_reference_topk(torch, rows_fp16, q_fp16, k, row_offset=0, chunk=1_000_000):
    """Checker: top-k of an fp32 torch scan of the same fp16 operands, in row chunks (never the [Q, N] matrix)."""
    Q, N = q_fp16.shape[0], rows_fp16.shape[0]
    dev = rows_fp16.device
    best_s = torch.full((Q, k), -float("inf"), device=dev)
    best_i = torch.full((Q, k), -1, dtype=torch.int64, device=dev)
    for r0 in range(0, N, chunk):
        s = q_fp16.float() @ rows_fp16[r0:r0 + chunk].float().t()
        cs, ci = torch.cat([best_s, s], 1).topk(min(k, best_s.shape[1] + s.shape[1]), dim=1)
        all_i = torch.cat([best_i, torch.arange(r0, r0 + s.shape[1], device=dev).expand(Q, -1) + row_offset], 1)
        best_s, best_i = cs, all_i.gather(1, ci)
    return best_s, best_i


def lists_agree(torch, s_a, i_a, s_b, i_b, eps):
    """Ranked lists equal up to ties: scores within eps position by position; where ids differ, the two rows involved
    score within eps of each other (a swap among near-equal scores or an exchange at the k-th boundary)."""
    if s_a.shape != s_b.shape or not bool(((s_a - s_b).abs() <= eps).all()):
        return False
    diff = i_a != i_b
    if not bool(diff.any()):
        return True
    # a differing id must appear in the other list at a position whose score is within eps, or sit at the boundary
    kth = torch.minimum(s_a[:, -1], s_b[:, -1]).unsqueeze(1)
    ok = torch.ones_like(diff)
    for q, p in diff.nonzero().tolist():
        where = (i_b[q] == i_a[q, p]).nonzero()
        if where.numel():
            ok[q, p] = bool((s_b[q, where[0, 0]] - s_b[q, p]).abs() <= eps)
        else:
            ok[q, p] = bool((s_a[q, p] - kth[q, 0]).abs() <= 2 * eps)
    return bool(ok.all())


def timeline_breakdown(timeline, batches):
    """Device-stamped (family, start, end) of one step -> ms per batch: busy time per family, the time no stamped kernel
    was running (launch gaps + the few unstamped small kernels), and how much neighbours overlapped (PDL)."""
    if not timeline:
        return None
    busy, overlap, idle = {}, 0, 0
    ordered = sorted(timeline, key=lambda t: t[1])
    frontier = ordered[0][1]
    for fam, t0, t1 in ordered:
        busy[fam] = busy.get(fam, 0) + (t1 - t0)
        if t0 > frontier:
            idle += t0 - frontier
        else:
            overlap += min(frontier, t1) - t0
        frontier = max(frontier, t1)
    span = frontier - ordered[0][1]
    out = {f: round(v * 1e-6 / batches, 4) for f, v in busy.items()}
    out.update(idle_between_kernels=round(idle * 1e-6 / batches, 4), overlapped=round(overlap * 1e-6 / batches, 4),
               span=round(span * 1e-6 / batches, 4), launches_per_batch=round(len(ordered) / batches, 1),
               avg_us={f: round(busy[f] * 1e-3 / sum(1 for t in ordered if t[0] == f), 2) for f in busy})
    return out


# ----------------------------------------------------------------------------------------------- secondary legs
def bench_encode_other(pkg, dev, name, batch, steps, warmup, peaks):
    """Secondary line: another tower geometry (BASELINE config 5 is ViT-L/14) on rank 0's GPU, device-resident frames."""
    import torch
    sd = pkg.clip.random_visual_state_dict(name, seed=1)
    enc = pkg.vit.VisionEncoder.from_state_dict(sd, name, device=dev, max_batch=batch, prefix="visual.")
    g = torch.Generator(device="cpu").manual_seed(7)
    frames = [torch.randint(0, 256, (batch, 224, 224, 3), generator=g, dtype=torch.uint8).to(dev) for _ in range(4)]
    for i in range(warmup):
        enc.encode(frames[i % 4], normalise=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for i in range(steps):
        enc.encode(frames[i % 4], normalise=True)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    total_flops, gemm_flops = vit_flops_per_frame(name)
    tf = total_flops * batch / (ms * 1e-3) / 1e12
    # in-situ split of four batches: device stamps of the GEMM / LayerNorm / attention launches, no host events
    pkg._lib.profile_enable(2)
    for i in range(4):
        enc.encode(frames[i % 4], normalise=True)
    torch.cuda.synchronize()
    in_situ = timeline_breakdown(pkg._lib.profile_timeline(), 4)
    pkg._lib.profile_enable(False)
    gemm_frac = None
    if in_situ and in_situ.get("gemm_tc"):
        gemm_frac = round(gemm_flops * batch / (in_situ["gemm_tc"] * 1e-3) / 1e12 / peaks["tf_sustained"], 4)
    return {"metric": f"frames/s {name} encode", "in_situ_ms_per_batch": in_situ, "gemm_frac_in_situ": gemm_frac, "unit": "frames/s", "value": round(batch / ms * 1e3, 1),
            "ms_per_step": round(ms, 4), "steps": steps, "warmup": warmup, "n_gpus": 1,
            "config": {"workload": f"{name} keyframe encoding, batch {batch} synthetic 224x224 frames, bf16, one GPU"},
            "whole_step_frac": round(tf / peaks["tf_sustained"], 4), "achieved_tflops": round(tf, 1),
            "timed_region_s": round(ms * steps * 1e-3, 3)}


def bench_config5(pkg, dev, peaks, n_frames=3600, n_caption_images=144, caps=5, batch=128):
    """BASELINE config 5 end to end on one GPU: ViT-L/14 (768-d) encode of a synthetic 1-hour 1-fps video set ->
    HBM store -> 720 caption queries -> R@1/5/10 + MRR by the reference harness's definitions
    (compare_models.py:999-1090, `metrics.evaluate_retrieval`), timed from pinned host frames to metrics.
    Checker (outside the timed region): the same metrics from an fp32 stock-PyTorch tower with the same weights
    (scripts/torch_eager_baseline.EagerViT) on the same frames and captions."""
    import numpy as np
    import torch
    name = "ViT-L/14"
    sd = pkg.clip.random_visual_state_dict(name, seed=1)
    enc = pkg.vit.VisionEncoder.from_state_dict(sd, name, device=dev, max_batch=batch, prefix="visual.")
    g = torch.Generator(device="cpu").manual_seed(55)
    # frames with low-frequency structure + noise: distinct embeddings (uniform noise frames all encode alike)
    frames = torch.empty((n_frames, 224, 224, 3), dtype=torch.uint8).pin_memory()
    for i in range(0, n_frames, 400):
        n = min(400, n_frames - i)
        low = torch.randint(0, 256, (n, 8, 8, 3), generator=g, dtype=torch.uint8)
        up = torch.nn.functional.interpolate(low.permute(0, 3, 1, 2).float(), size=224, mode="bilinear")
        noise = torch.randint(0, 64, (n, 3, 224, 224), generator=g, dtype=torch.uint8).float()
        frames[i:i + n] = (0.75 * up + noise).clamp(0, 255).to(torch.uint8).permute(0, 2, 3, 1)
    host_batches = [frames[i:i + batch] for i in range(0, n_frames, batch)]
    with_caps = torch.randperm(n_frames, generator=g)[:n_caption_images].sort().values.numpy()
    caption_image = np.repeat(with_caps, caps)
    cap_noise = torch.randn((len(caption_image), 768), generator=g).numpy().astype(np.float32)

    def captions(image_emb):   # captions = noisy copies of their image's embedding (5 per sampled image)
        c = image_emb[caption_image] + 0.6 * cap_noise / np.sqrt(768)
        return c / np.linalg.norm(c, axis=-1, keepdims=True)

    for _ in enc.encode_stream(host_batches[:3], normalise=True):   # warm-up (3 batches)
        pass
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    emb = torch.cat(list(enc.encode_stream(host_batches, normalise=True))).numpy()
    t_encode = time.perf_counter() - t0
    # the checker's captions are made from ITS embeddings below; the timed path uses captions of its own embeddings
    text = captions(emb)
    res = pkg.metrics.evaluate_retrieval(emb, text, caption_image, device=dev)
    torch.cuda.synchronize()
    t_total = time.perf_counter() - t0
    total_flops, _ = vit_flops_per_frame(name)

    eager = eager_module()
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    ref_model = eager.EagerViT(patch=14, d=1024, layers=24, heads=16, out=768).to(dev).float().eval()
    ref_model.load_openai_state_dict({k: v.to(dev) for k, v in sd.items()})
    with torch.no_grad():
        ref_emb = torch.cat([ref_model(frames[i:i + 64].to(dev)) for i in range(0, n_frames, 64)]).cpu().numpy()
    del ref_model
    torch.cuda.empty_cache()
    cos = (emb * ref_emb).sum(-1)
    # same captions for both stores (made from the fp32 embeddings): the metrics may differ only through the tower
    text_ref = captions(ref_emb)
    ours_on_ref_caps = pkg.metrics.evaluate_retrieval(emb, text_ref, caption_image, device=dev)
    sim = ref_emb @ text_ref.T                                           # the reference's own numpy formulation
    t2i = np.array([int(np.where(np.argsort(-sim[:, i]) == caption_image[i])[0][0]) + 1
                    for i in range(sim.shape[1])])
    ref_t2i = pkg.metrics.metrics_from_ranks(t2i)
    keys = ("R@1", "R@5", "R@10", "MRR")
    delta = {k: round(abs(ours_on_ref_caps["t2i"][k] - ref_t2i[k]), 5) for k in keys}
    return {
        "metric": "ViT-L/14 encode + index + query, R@K / MRR (BASELINE config 5)", "n_gpus": 1,
        "config": {"workload": f"{n_frames} synthetic 224x224 frames (1 h at 1 fps) -> ViT-L/14 -> 768-d fp16 store -> "
                               f"{len(caption_image)} caption queries ({caps} noisy copies of {n_caption_images} image "
                               "embeddings), full rankings through the fused scan"},
        "seconds_end_to_end": round(t_total, 3), "seconds_encode": round(t_encode, 3),
        "frames_per_s_end_to_end": round(n_frames / t_total, 1),
        "encode_frac_of_sustained_peak": round(total_flops * n_frames / t_encode / 1e12 / peaks["tf_sustained"], 4),
        "metrics": {d: {k: round(v, 5) for k, v in res[d].items()} for d in ("t2i", "i2t", "mean")},
        "parity": {"checker": "fp32 stock-PyTorch ViT-L/14 with the same weights (TF32 off), numpy argsort ranks",
                   "min_cosine": round(float(cos.min()), 6), "cosine_ok": bool(cos.min() >= 0.999),
                   "t2i_checker": {k: round(ref_t2i[k], 5) for k in keys},
                   "t2i_ours_same_captions": {k: round(ours_on_ref_caps["t2i"][k], 5) for k in keys},
                   "abs_delta": delta,
                   "tolerance": "R@K within max(2 captions, 0.01), MRR within 0.01 (tests/test_vit_gpu.py: random-init "
                                "towers give near-identical embeddings, ranks move where two cosines tie within the "
                                "bf16 tower's error)",
                   "ok": bool(cos.min() >= 0.999 and delta["MRR"] <= 0.01 and
                              all(delta[k] <= max(2.0 / len(caption_image), 0.01) + 1e-9 for k in keys[:3]))}}


def bench_scenecut(pkg, dev, peaks, frames_n=256, steps=10, warmup=3):
    """Secondary line: scene-cut scoring of 1280x720 frames resident in HBM (707 MB > L2)."""
    import importlib

    import torch
    scenecut = importlib.import_module(pkg.__name__ + ".scenecut")
    H, W = 720, 1280
    g = torch.Generator(device=dev).manual_seed(11)
    frames = torch.randint(0, 256, (frames_n, H, W, 3), dtype=torch.uint8, device=dev, generator=g)
    det = scenecut.ContentDetector(device=dev)
    for _ in range(warmup):
        det.delta_sums(frames)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        det.delta_sums(frames)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    gbs = frames_n * H * W * 3 / (ms * 1e-3) / 1e9
    return {"metric": "frames/s scene-cut scoring (HSV frame difference)", "unit": "frames/s",
            "value": round(frames_n / ms * 1e3, 1), "ms_per_step": round(ms, 4), "steps": steps, "warmup": warmup,
            "n_gpus": 1, "dtype": "u8/int32",
            "config": {"workload": f"{frames_n} consecutive 1280x720 BGR frames, device-resident, per-pair |dH|,|dS|,|dV| sums"},
            "roofline": {"bound": "hbm", "achieved": round(gbs, 1), "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": round(gbs / peaks["hbm_gbs"], 4), "traffic": None,
                         "note": "algorithmic bytes = every frame once (3 B per pixel); the kernel is bound by the "
                                 "integer pipe on the exact OpenCV HSV arithmetic (profiles/r01_e_ncu_scenecut.txt)"}}


def bench_ingest_jpeg(pkg, dev, peaks, n_frames=4096, pil_sample=512):
    """Secondary line: folder ingest end to end -- 1280x720 JPEG files -> nvJPEG batched decode ->
    resize + crop -> ViT-B/32 -> `.npy`, through `EmbeddingService.extract_and_save_embeddings_from_folder(gpu_jpeg=True)`;
    the same call with PIL on the host threads (the exact mode) on a bounded sample beside it."""
    import io
    import shutil
    import tempfile

    import numpy as np
    fixtures = os.path.join(ROOT, "tests", "golden", "ref_frames")
    blobs = [open(os.path.join(fixtures, f), "rb").read() for f in sorted(os.listdir(fixtures)) if f.endswith(".jpg")] \
        if os.path.isdir(fixtures) else []
    source = f"the reference's own key-frames ({len(blobs)} files from tests/golden/ref_frames, repeated)"
    if not blobs:   # synthetic stand-ins: smooth colour fields + grain, cv2.imwrite's default quality
        from PIL import Image
        rng = np.random.default_rng(5)
        yy, xx = np.mgrid[0:720, 0:1280].astype(np.float32)
        for i in range(8):
            f = [rng.uniform(0.002, 0.02, 2) for _ in range(3)]
            img = np.stack([127 + 90 * np.sin(a * xx + i) * np.cos(b * yy) for a, b in f], -1)
            img = np.clip(img + rng.normal(0, 6, img.shape), 0, 255).astype(np.uint8)
            buf = io.BytesIO()
            Image.fromarray(img).save(buf, "JPEG", quality=95)
            blobs.append(buf.getvalue())
        source = "8 synthetic 1280x720 JPEGs (quality 95), repeated"
    tmp = tempfile.mkdtemp(prefix="evr_bench_ingest_")
    try:
        for sub in ("frames/vid", "frames/pil", "metadata", "embedding"):
            os.makedirs(os.path.join(tmp, sub))
        nbytes = 0
        for i in range(n_frames):
            with open(os.path.join(tmp, "frames", "vid", f"{i}.jpg"), "wb") as f:
                f.write(blobs[i % len(blobs)])
            nbytes += len(blobs[i % len(blobs)])
            if i < pil_sample:
                with open(os.path.join(tmp, "frames", "pil", f"{i}.jpg"), "wb") as f:
                    f.write(blobs[i % len(blobs)])
        sd = pkg.clip.random_visual_state_dict("ViT-B/32", seed=1)
        out, rows = {}, {}
        for mode, folder, n in ((True, "vid", n_frames), (False, "pil", pil_sample)):
            ps = pkg.services.PathService(os.path.join(tmp, "frames"), os.path.join(tmp, "metadata"),
                                          os.path.join(tmp, "embedding"))
            cs = pkg.services.CacheService()
            with contextlib.redirect_stdout(io.StringIO()):   # the service prints like the reference's does
                svc = pkg.services.EmbeddingService(cs, ps, pkg.services.DataService(ps, cs), dev,
                                                    load_kwargs=dict(state_dict=sd, max_batch=256), gpu_jpeg=mode)
                path = os.path.join(tmp, "frames", folder)
                svc.extract_and_save_embeddings_from_folder(path, None, "warm")     # pinned buffers, plans, nvJPEG state
                launches0 = pkg._lib.launch_count()
                t0 = time.perf_counter()
                npy = svc.extract_and_save_embeddings_from_folder(path, None, folder)
                dt = time.perf_counter() - t0
            rows[mode] = np.load(npy).astype(np.float32)
            out[mode] = {"frames_per_s": round(n / dt, 1), "seconds": round(dt, 3), "frames": n,
                         "gpu_launches": pkg._lib.launch_count() - launches0}
            if mode:
                out[mode]["backend"] = svc._jpeg.backend
        # rows follow sorted(os.listdir(folder)) (embedding_service.py:452): align the two folders by file name
        where = {name: i for i, name in enumerate(sorted(f"{i}.jpg" for i in range(n_frames)))}
        a = rows[False]
        b = rows[True][[where[name] for name in sorted(f"{i}.jpg" for i in range(pil_sample))]]
        cos = (a * b).sum(-1) / (np.linalg.norm(a, axis=-1) * np.linalg.norm(b, axis=-1))
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    fps = out[True]["frames_per_s"]
    moved = nbytes / n_frames + 2 * 720 * 1280 * 3 + 2 * 224 * 224 * 3    # JPEG in, RGB written + read, crop written + read
    gbs = fps * moved / 1e9
    return {"metric": "frames/s folder ingest, 1280x720 JPEG files -> .npy rows (ViT-B/32)", "unit": "frames/s",
            "value": fps, "n_gpus": 1, "dtype": "u8 -> bf16", "seconds": out[True]["seconds"],
            "gpu_launches": out[True]["gpu_launches"],
            "config": {"workload": f"{n_frames} files, {nbytes / n_frames / 1024:.1f} KB mean; {source}; wall clock of the "
                                   "reference-facing call, files in the page cache, random-init weights",
                       "decode": "nvJPEG " + out[True].get("backend", "?") + ", 256-frame batches, 3 states in flight"},
            "pil_threads": out[False],
            "parity": {"min_cosine_vs_pil_path": round(float(cos.min()), 6), "rows": int(len(cos)), "tolerance": 0.999,
                       "ok": bool(cos.min() >= 0.999)},
            "roofline": {"bound": "hbm", "achieved": round(gbs, 1), "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": round(gbs / peaks["hbm_gbs"], 4), "traffic": None,
                         "note": "algorithmic bytes per frame = JPEG bytes + decoded RGB written and read once + crop "
                                 "written and read once; the line is bound by nvJPEG's Huffman stage (a library: "
                                 "~20 ms per batched call whatever the batch), not by HBM -- the B200's hardware JPEG "
                                 "engines are not reachable through CUDA 12.9's nvJPEG (status 7 at create)"}}


def bench_library(dev, batch, k, compile_too):
    """`library_baseline`: stock PyTorch on the same B200 (SURVEY.md 2.1's second baseline) -- cuBLAS + SDPA + ATen
    LayerNorm in bf16 for the tower, `matmul` + `topk` in 500 k-row chunks for the scan.  Not the product."""
    import argparse as ap

    import torch
    eager = eager_module()
    out = {"torch": torch.__version__}
    a = ap.Namespace(model=VIT_NAME, batch=batch, steps=20, rows=10_000_000, queries=10_000, chunk=500_000)
    r = eager.encode(a)
    out["encode_eager"] = {"value": round(r["value"], 1), "unit": "frames/s", "ms_per_batch": round(r["ms_per_step"], 4)}
    if compile_too:
        try:
            t0 = time.perf_counter()
            m = eager.EagerViT().to(dev).to(torch.bfloat16).eval()
            cm = torch.compile(m)
            gen = torch.Generator(device=dev).manual_seed(0)
            pool = [torch.randint(0, 256, (batch, 224, 224, 3), dtype=torch.uint8, device=dev, generator=gen)
                    for _ in range(8)]
            i = [0]

            def step():
                with torch.no_grad():
                    cm(pool[i[0] % 8])
                i[0] += 1
            ms = eager.timed(step, 20)
            out["encode_compiled"] = {"value": round(batch / ms * 1e3, 1), "unit": "frames/s",
                                      "ms_per_batch": round(ms, 4),
                                      "compile_and_run_s": round(time.perf_counter() - t0, 1)}
            del cm, m, pool
        except Exception as e:  # inductor / triton unavailable on the box
            out["encode_compiled"] = {"unavailable": f"{type(e).__name__}: {str(e)[:160]}"}
    torch.cuda.empty_cache()
    a.steps = 3
    r = eager.search(a)
    out["search_eager"] = {"value": round(r["value"], 1), "unit": "queries/s", "ms_per_step": round(r["ms_per_step"], 2),
                           "k": k}
    torch.cuda.empty_cache()
    return out


# ----------------------------------------------------------------------------------------------- encode (ours)
def bench_encode(pkg, dev, rank, world, steps, warmup, batch, inner, peaks):
    import numpy as np
    import torch
    _lib = pkg._lib
    # seeded random init with openai's scales (no weights on the box); the product's own initialiser, not the oracle's
    sd = pkg.clip.random_visual_state_dict(VIT_NAME, seed=1)
    enc = pkg.vit.VisionEncoder.from_state_dict(sd, VIT_NAME, device=dev, max_batch=batch, prefix="visual.")
    # synthetic uint8 RGB frames; R distinct batches so consecutive batches never reuse L2 lines
    R = 8

    def rank_batches(r, count=R):
        g = torch.Generator(device="cpu").manual_seed(1000 + r)
        return [torch.randint(0, 256, (batch, 224, 224, 3), generator=g, dtype=torch.uint8) for _ in range(count)]

    host = [h.pin_memory() for h in rank_batches(rank)]
    resident = [h.to(dev) for h in host]
    out = None

    def one_step(s):
        nonlocal out
        for j in range(inner):
            out = enc.encode(resident[(s * inner + j) % R], normalise=True)

    for s in range(warmup):
        one_step(s)
    barrier_sync(world)
    sampler = ClockSampler(dev.index)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    launches0 = _lib.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier_sync(world)
    e0.record()
    for s in range(steps):
        one_step(s)
    e1.record()
    barrier_sync(world)
    ms = max_over_ranks(e0.elapsed_time(e1), dev, world)
    launches = _lib.launch_count() - launches0
    if rank == 0:
        time.sleep(0.3)
    clocks = sampler.stop() if rank == 0 else None
    assert torch.isfinite(out).all()

    # ---- end to end through the public host API, H2D + D2H inside the timed region
    e2e_steps = steps if world == 1 else min(steps, 5)
    n_local = e2e_steps * inner * batch
    for _ in enc.encode_stream((host[i % R] for i in range(8)), normalise=True):
        pass
    barrier_sync(world)
    parity = {}
    if world == 1:
        t0 = time.perf_counter()
        e0.record()
        n_out = 0
        for emb in enc.encode_stream((host[i % R] for i in range(e2e_steps * inner)), normalise=True):
            n_out += emb.shape[0]
        e1.record()
        barrier_sync(world)
        wall_ms = (time.perf_counter() - t0) * 1e3
        e2e_ms = max(e0.elapsed_time(e1), wall_ms)
        assert n_out == n_local
        e2e_api = "VisionEncoder.encode_stream(pinned uint8 [256,224,224,3] batches) -> host fp32 [256,512] rows"
        d2h = batch * 512 * 4 * inner
    else:
        # data-parallel ingest: the global frame list is the rank slices end to end; every rank encodes its
        # contiguous slice and writes its rows into ONE pre-sized .npy (services.encode_frames_sharded)
        import torch.distributed as dist
        n_total = n_local * world
        lo, hi = pkg.search.shard_bounds(n_total, world, rank)
        assert hi - lo == n_local
        path_box = [os.path.join(tempfile.gettempdir(), f"evr_bench_dp_{os.environ.get('MASTER_PORT', '0')}.npy")]
        t0 = time.perf_counter()
        e0.record()
        rows = pkg.services.encode_frames_sharded(enc, (host[i % R] for i in range(e2e_steps * inner)), n_total, lo,
                                                  path_box[0], rank, world, dist.barrier)
        e1.record()
        barrier_sync(world)
        wall_ms = (time.perf_counter() - t0) * 1e3
        e2e_ms = max_over_ranks(max(e0.elapsed_time(e1), wall_ms), dev, world)
        e2e_api = ("services.encode_frames_sharded: rank slices of one frame list -> VisionEncoder.encode_stream -> "
                   "one shared .npy (fp16 unit-norm rows)")
        d2h = batch * 512 * 2 * inner
        # parity of the shared file (outside the timed region): rank 0 re-reads rows of EVERY rank's slice and
        # compares them bit for bit with its own encode of the frames that rank drew (same seeds)
        if rank == 0:
            mm = np.load(path_box[0], mmap_mode="r")
            ok, checked = mm.shape == (n_total, 512), 0
            for r in range(world):
                theirs = rank_batches(r, 1)[0][:16]
                mine = enc.encode(theirs.to(dev), normalise=True, out_dtype=torch.float16).cpu().numpy()
                r_lo = pkg.search.shard_bounds(n_total, world, r)[0]
                ok = ok and np.array_equal(np.asarray(mm[r_lo:r_lo + 16]), mine)
                ok = ok and np.array_equal(np.asarray(mm[r_lo + R * batch:r_lo + R * batch + 16]), mine)  # batch R = batch 0
                checked += 32
            parity["dp_ingest_file"] = {"rows_total": int(n_total), "rows_checked": checked, "bit_identical": bool(ok),
                                        "what": "rows of every rank's slice of the shared .npy == rank 0's own encode of "
                                                "the same frames"}
            del mm
        barrier_sync(world)
        if rank == 0:
            os.unlink(path_box[0])

    # ---- parity of the timed configuration (outside the timed regions): 4 frames vs an fp32 stock-PyTorch tower
    if rank == 0:
        eager = eager_module()
        torch.backends.cuda.matmul.allow_tf32 = False
        torch.backends.cudnn.allow_tf32 = False
        ref_model = eager.EagerViT().to(dev).float().eval()
        ref_model.load_openai_state_dict({k: v.to(dev) for k, v in sd.items()})
        with torch.no_grad():
            ref = ref_model(resident[0][:4])
        ours = enc.encode(resident[0], normalise=True)[:4]          # the timed batch size
        cos = (ours * ref).sum(-1)
        single = enc.encode(resident[0][:4], normalise=True)
        parity["encode"] = {"frames": 4, "min_cosine_vs_fp32_tower": round(float(cos.min()), 6),
                            "ok": bool(cos.min() >= 0.999), "tolerance": "cosine >= 0.999 per row (north_star)",
                            "batch_invariant": bool(torch.equal(single, ours)),
                            "checker": "scripts/torch_eager_baseline.EagerViT in fp32 (TF32 off) with the same weights"}
        del ref_model

    # ---- in-situ kernel time (separate passes of ONE step = `inner` batches)
    roof = None
    if rank == 0:
        total_flops, gemm_flops = vit_flops_per_frame(VIT_NAME)
        # pass A, no events: every GEMM launch stamps %globaltimer on the device (first CTA past
        # griddepcontrol.wait -> last CTA exit), programmatic overlap intact: the in-situ figure
        _lib.profile_enable(2)
        one_step(0)
        torch.cuda.synchronize()
        dt_ms, dt_n = _lib.profile_read()["gemm_tc"]
        in_situ = timeline_breakdown(_lib.profile_timeline(), inner)
        # pass B, CUDA events around every launch: per-family split (serialises the overlap, +~5 us per launch)
        _lib.profile_enable(True)
        one_step(0)
        torch.cuda.synchronize()
        fam = _lib.profile_read()
        _lib.profile_enable(False)
        gemm_ms, gemm_n = fam["gemm_tc"]
        all_ms = sum(v[0] for v in fam.values())
        step_ms = ms / steps
        ach = gemm_flops * batch * inner / (dt_ms * 1e-3) / 1e12 if dt_n else None
        ev_ach = gemm_flops * batch * inner / (gemm_ms * 1e-3) / 1e12
        whole = total_flops * batch * inner / (step_ms * 1e-3) / 1e12
        roof = {
            "bound": "tensor", "kernel": "gemm_tc_kernel (tcgen05, all ViT GEMMs)",
            "achieved": round(ach if ach else ev_ach, 1), "peak": peaks["tf_sustained"], "unit": "TFLOP/s",
            "frac": round((ach if ach else ev_ach) / peaks["tf_sustained"], 4),
            "traffic": ncu_traffic("gemm_tc_kernel"),
            "peak_source": peaks["source"] + ", sustained bf16 (the timed region is seconds long)",
            "timing": "in situ: each GEMM launch stamps %globaltimer on the device (first CTA past griddepcontrol.wait "
                      "-> last CTA exit) during a pass of one step without host events"
                      if ach else "CUDA events around every launch",
            "flops_per_launch_avg": gemm_flops * batch * inner / max(dt_n or gemm_n, 1),
            "avg_launch_ms": round((dt_ms / dt_n) if dt_n else gemm_ms / max(gemm_n, 1), 5),
            "launches_per_step": dt_n or gemm_n,
            "gemm_ms_per_batch": round((dt_ms if dt_n else gemm_ms) / inner, 4),
            "share_of_step": round((dt_ms if dt_n else gemm_ms) / step_ms, 4),
            "whole_step_frac": round(whole / peaks["tf_sustained"], 4), "whole_step_tflops": round(whole, 1),
            "whole_step_frac_of_burst_peak": round(whole / peaks["tf_burst"], 4),
            "in_situ_ms_per_batch": in_situ,
            "event_timed": {"achieved": round(ev_ach, 1), "frac": round(ev_ach / peaks["tf_sustained"], 4),
                            "family_ms_per_batch": {k: round(v[0] / inner, 4) for k, v in fam.items() if v[1]},
                            "share_of_step": round(gemm_ms / all_ms, 4),
                            "note": "CUDA events around every launch serialise the PDL overlap (+~5 us per launch): "
                                    "use for the split between families, not for the roofline fraction"},
        }
    frames = batch * inner * world * steps
    e2e_frames = batch * inner * world * e2e_steps
    return {
        "value": frames / (ms * 1e-3), "ms_per_step": ms / steps, "gpu_launches": int(launches), "clocks": clocks,
        "timed_region_s": ms * 1e-3,
        "e2e": {"value": round(e2e_frames / (e2e_ms * 1e-3), 1), "unit": "frames/s",
                "h2d_bytes_per_step": batch * 224 * 224 * 3 * inner, "d2h_bytes_per_step": d2h,
                "steps": e2e_steps, "timed_region_s": round(e2e_ms * 1e-3, 3), "api": e2e_api},
        "roofline": roof, "parity": parity,
    }


# ----------------------------------------------------------------------------------------------- search (ours)
def make_store(torch, dev, n_rows, seed):
    """[n_rows, 512] unit-norm fp16 rows generated on the device (N(0,1) -> L2 normalise -> fp16)."""
    g = torch.Generator(device=dev).manual_seed(seed)
    rows = torch.empty((n_rows, 512), dtype=torch.float16, device=dev)
    chunk = 1_000_000
    for r0 in range(0, n_rows, chunk):
        n = min(chunk, n_rows - r0)
        x = torch.randn((n, 512), generator=g, device=dev)
        rows[r0:r0 + n] = (x / x.norm(dim=-1, keepdim=True)).half()
    return rows


def time_steps(torch, fn, steps, warmup, world):
    for _ in range(warmup):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier_sync(world)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    barrier_sync(world)
    return e0.elapsed_time(e1)


def search_parity(pkg, torch, dev, rank, world, store, sharded, q, k, total_rows, lo, hi, planted, n_check=32):
    """Checks of the timed search configuration's own results (outside the timed region):
    planted rows come first; `n_check` sampled queries agree with a chunked fp32 torch scan of the same operands
    (local shard, then -- through an all-gather of the checker's lists -- the merged global list); the merged lists
    are bit-identical on every rank."""
    import torch.distributed as dist
    eps = 1e-5   # same fp16 operands, fp32 accumulation in a different order
    sc, idx = sharded.search(q, k)
    P = planted.numel()
    out = {"planted_rows_first": bool(torch.equal(idx[:P, 0].cpu(), planted.cpu())),
           "planted_score_is_1": bool(((sc[:P, 0] - 1.0).abs() < 2e-3).all()),
           "sorted_desc": bool((sc[:, 1:] <= sc[:, :-1]).all()),
           "ids_in_range": bool(int(idx.min()) >= 0 and int(idx.max()) < total_rows)}
    srt = idx.sort(dim=1).values
    out["no_duplicates"] = bool((srt[:, 1:] != srt[:, :-1]).all())
    g = torch.Generator(device="cpu").manual_seed(99)
    pick = torch.randperm(q.shape[0], generator=g)[:n_check].to(dev)
    ref_s, ref_i = chunked_reference_topk(torch, store.rows, q[pick], k, row_offset=lo)
    loc_s, loc_i = store.search(q, k)
    out["local_vs_fp32_scan"] = {"queries": int(pick.numel()), "rows": int(hi - lo), "eps": eps,
                                 "ok": lists_agree(torch, loc_s[pick], loc_i[pick], ref_s, ref_i, eps)}
    if world > 1:
        all_s = torch.empty((world,) + tuple(ref_s.shape), dtype=ref_s.dtype, device=dev)
        all_i = torch.empty((world,) + tuple(ref_i.shape), dtype=ref_i.dtype, device=dev)
        dist.all_gather_into_tensor(all_s.view(-1, k), ref_s.contiguous())
        dist.all_gather_into_tensor(all_i.view(-1, k), ref_i.contiguous())
        cat_s = all_s.permute(1, 0, 2).reshape(pick.numel(), world * k)
        cat_i = all_i.permute(1, 0, 2).reshape(pick.numel(), world * k)
        top_s, sel = cat_s.topk(k, dim=1)
        top_i = cat_i.gather(1, sel)
        out["merged_vs_fp32_scan"] = {"queries": int(pick.numel()), "rows": int(total_rows), "eps": eps,
                                      "ok": lists_agree(torch, sc[pick], idx[pick], top_s, top_i, eps)}
        # bit-identical merged lists on every rank: a position-weighted hash, min over ranks == max over ranks
        w = torch.arange(1, idx.numel() + 1, device=dev, dtype=torch.int64).view_as(idx) % 1_000_003
        h = float(((idx * w) % 2_147_483_647).sum().item() % 2_147_483_647) + \
            float((sc.view(torch.int32).to(torch.int64) * w % 2_147_483_647).sum().item() % 2_147_483_647)
        out["merged_identical_on_all_ranks"] = bool(reduce_over_ranks(h, dev, world, "min") ==
                                                    reduce_over_ranks(h, dev, world, "max"))
    flags = [v if isinstance(v, bool) else v["ok"] for v in out.values()]
    ok = float(all(flags))
    out["ok"] = bool(reduce_over_ranks(ok, dev, world, "min") == 1.0)
    return out


def bench_search(pkg, dev, rank, world, steps, warmup, total_rows, Q, k, peaks, hbm_regime=True, label=None):
    import torch
    _lib = pkg._lib
    D = 512
    lo, hi = pkg.search.shard_bounds(total_rows, world, rank)
    rows = make_store(torch, dev, hi - lo, seed=2 + rank)
    g = torch.Generator(device="cpu").manual_seed(3)
    qh = torch.randn((Q, D), generator=g)
    qh = (qh / qh.norm(dim=-1, keepdim=True))
    q = qh.to(dev).half()
    qh = q.float().cpu().pin_memory()        # host queries = exactly the fp16 values (so planted rows score 1.0)
    # planted rows: query j's own vector stored at a known global row, spread over all shards
    P = min(64, Q)
    planted = (torch.arange(P, dtype=torch.int64) * (total_rows // P) + 17) % total_rows
    mine = (planted >= lo) & (planted < hi)
    rows[(planted[mine] - lo).to(dev)] = q[:P][mine.to(dev)]
    store = pkg.search.FrameStore(rows, row_offset=lo)
    sharded = pkg.search.ShardedFrameStore(store) if world > 1 else store
    res = {}
    n_local = hi - lo
    # ---- HBM-bound regime first: small query batches stream the whole local shard once per step.  Measured before
    # the tensor-bound batch so each regime is timed in its own steady state (a serving GPU answering single queries
    # is not sitting at the power cap a 10 000-query batch drives it to).
    regime = []
    if hbm_regime:
        for qs in (1, 8, 32, 128):
            qq = q[:qs].contiguous()
            t = max_over_ranks(time_steps(torch, lambda: sharded.search(qq, k), max(steps, 10), 3, world), dev,
                               world) / max(steps, 10)
            b = n_local * D * 2 + qs * D * 2 + qs * k * 12
            gbs = b / (t * 1e-3) / 1e9
            regime.append({"Q": qs, "ms": round(t, 4), "queries_per_s": round(qs / (t * 1e-3), 1),
                           "GB_per_s_per_gpu": round(gbs, 1), "frac_of_hbm_peak": round(gbs / peaks["hbm_gbs"], 4)})

    def step():
        res["out"] = sharded.search(q, k)

    launches0 = _lib.launch_count()
    ms = max_over_ranks(time_steps(torch, step, steps, warmup, world), dev, world)
    launches = (_lib.launch_count() - launches0) * steps // (steps + warmup)

    # ---- end to end: pinned fp32 queries on the host -> H2D -> scan -> D2H of (scores, ids)
    def step_e2e():
        s, i = sharded.search(qh.to(dev, non_blocking=True), k)
        res["host"] = (s.cpu(), i.cpu())

    e2e_ms = max_over_ranks(time_steps(torch, step_e2e, steps, 2, world), dev, world)
    parity = search_parity(pkg, torch, dev, rank, world, store, sharded, q, k, total_rows, lo, hi, planted)

    flops = 2.0 * Q * n_local * D
    nbytes = n_local * D * 2 + Q * D * 2 + Q * k * 12
    out = {
        "metric": label or (SEARCH_METRIC if world == 1 else "queries/s top-100 over 100M-frame store row-sharded"),
        "unit": "queries/s", "value": round(Q * steps / (ms * 1e-3), 1), "ms_per_step": round(ms / steps, 3),
        "steps": steps, "warmup": warmup, "gpu_launches": int(launches),
        "scaling": "strong" if total_rows >= 100_000_000 else "weak", "timed_region_s": round(ms * 1e-3, 3),
        "config": {"workload": f"cosine top-{k}, {Q} queries x {total_rows} frames x {D}-d fp16 store"
                               + (f", row-sharded over {world} GPUs + NCCL all-gather + merge" if world > 1 else ""),
                   "l2": "store (>= 10 GB per GPU) is far larger than L2"},
        "e2e": {"value": round(Q * steps / (e2e_ms * 1e-3), 1), "unit": "queries/s",
                "h2d_bytes_per_step": Q * D * 4, "d2h_bytes_per_step": Q * k * 12,
                "api": "FrameStore.search(pinned fp32 queries) -> host (scores, ids)"},
        "parity": parity,
    }
    if rank == 0:
        # dominant kernel, timed live (separate pass): algorithmic FLOPs (and bytes) per launch / duration
        _lib.profile_enable(True)
        for _ in range(steps):
            store.search(q, k)
        torch.cuda.synchronize()
        fam = _lib.profile_read()
        _lib.profile_enable(False)
        # a search call launches the scan kernel twice (sample pre-pass + full scan): time per call = family / steps
        scan_ms = fam["topk_scan"][0] / steps
        t_tc, t_hbm = flops / (peaks["tf_sustained"] * 1e12), nbytes / (peaks["hbm_gbs"] * 1e9)
        tensor_bound = t_tc >= t_hbm
        ach_tf = flops / (scan_ms * 1e-3) / 1e12
        ach_gb = nbytes / (scan_ms * 1e-3) / 1e9
        out["roofline"] = {
            "bound": "tensor" if tensor_bound else "hbm", "kernel": "topk_scan_kernel (tcgen05 GEMM + fused top-k)",
            "achieved": round(ach_tf if tensor_bound else ach_gb, 1),
            "peak": peaks["tf_sustained"] if tensor_bound else peaks["hbm_gbs"],
            "unit": "TFLOP/s" if tensor_bound else "GB/s",
            "frac": round((ach_tf / peaks["tf_sustained"]) if tensor_bound else (ach_gb / peaks["hbm_gbs"]), 4),
            "traffic": ncu_traffic("topk_scan_kernel") if total_rows == 10_000_000 and world == 1 else None,
            "algorithmic_bytes": nbytes, "peak_source": peaks["source"], "avg_launch_ms": round(scan_ms, 4),
            "launches_per_step": fam["topk_scan"][1] // steps, "achieved_gbs": round(ach_gb, 1), "achieved_tflops": round(ach_tf, 1),
            "share_of_step": round(fam["topk_scan"][0] / sum(v[0] for v in fam.values()), 4),
            "family_ms_per_step": {kk: round(v[0] / steps, 4) for kk, v in fam.items() if v[1]},
            "timing": "CUDA events around every launch (2 scan launches of >= 1 ms per call: event overhead < 0.1 %), "
                      "separate pass of the same K steps",
        }
    if hbm_regime:
        out["hbm_regime"] = regime
    del store, sharded, rows
    torch.cuda.empty_cache()
    return out


# ----------------------------------------------------------------------------------------------- CPU legs (oracle)
def cpu_encode(vo, n_batches, warm=1, batch=32):
    """The reference's ingest loop (embedding_service.py:467-502: batches of 32, no_grad, .cpu().numpy(), final
    L2 normalise) over its own fp32 CPU ViT, restated in oracle/vit_oracle.py.  Returns (frames/s, threads)."""
    import numpy as np
    import torch
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    visual = vo.build_visual(VIT_NAME, seed=1)
    _, pix = vo.synthetic_frames(batch * (n_batches + warm), seed=0)
    vo.encode_image(visual, pix[:batch * warm], batch=batch)
    t0 = time.perf_counter()
    e = vo.encode_image(visual, pix[batch * warm:], batch=batch)
    dt = time.perf_counter() - t0
    assert e.shape == (batch * n_batches, 512) and np.isfinite(e).all()
    return batch * n_batches / dt, threads


def cpu_search(so, n_rows, n_queries, k):
    """The reference's scan (embedding_service.py:314-320: np.dot + full argsort per query, no query batching) over an
    fp32 store of n_rows.  Returns queries/s at n_rows."""
    import numpy as np
    rng = np.random.default_rng(2)
    emb = so.normalise_store(rng.standard_normal((n_rows, 512), dtype=np.float32))
    q = rng.standard_normal((n_queries + 1, 512), dtype=np.float32)
    q /= np.linalg.norm(q, axis=-1, keepdims=True)
    so.search_top_frames(emb, q[:1], k)
    t0 = time.perf_counter()
    for j in range(1, n_queries + 1):
        so.search_top_frames(emb, q[j:j + 1], k)
    return n_queries / (time.perf_counter() - t0)


def encode_workload(args):
    return (f"ViT-B/32 keyframe encoding, batches of {args.batch} synthetic 224x224 frames per GPU, bf16 tensor cores, "
            "random-init weights (seeded)")


def run_reference(args):
    """--impl reference: the reference's CPU path for the same metric/config on all host threads (rank 0 only)."""
    rank, _, world = dist_env()
    if rank != 0:
        return
    from oracle import search_oracle as so, vit_oracle as vo
    import torch
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    visual = vo.build_visual(VIT_NAME, seed=1)
    batch = 32  # the reference's own batch size (embedding_service.py:461)
    _, pix = vo.synthetic_frames(batch, seed=0)
    for _ in range(max(1, min(args.warmup, 2))):
        vo.encode_image(visual, pix, batch=batch)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        vo.encode_image(visual, pix, batch=batch)
    dt = time.perf_counter() - t0
    fps = batch * args.steps / dt
    sample_rows = 1_000_000
    qps_1m = cpu_search(so, sample_rows, 4, args.k)
    line = {
        "impl": "reference", "metric": METRIC, "value": round(fps, 2), "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 2),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": encode_workload(args),
                   "step": "bounded sample: one batch of 32 frames (the reference's own batch size, "
                           "embedding_service.py:461) in fp32 on the host cores"},
        "cpu_baseline": {"value": round(fps, 2), "unit": "frames/s", "cores": threads, "kind": "port",
                         "sample": f"{args.steps} batches of 32 frames, oracle/vit_oracle.py (fp32 torch CPU restatement "
                                   "of openai/CLIP encode_image; the reference's `clip` submodule is absent)"},
        "e2e": {"value": round(fps, 2), "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "search": {"impl": "reference", "metric": SEARCH_METRIC, "unit": "queries/s",
                   "value": round(qps_1m * sample_rows / 10_000_000, 3),
                   "sample": f"4 queries, top-{args.k}, np.dot + full argsort over {sample_rows} fp32 rows "
                             f"({qps_1m:.2f} q/s), scaled linearly to 10 M rows (favours the reference: argsort is "
                             "N log N)", "cores": threads},
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------- main
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="both", choices=["both", "encode", "search"])
    ap.add_argument("--batch", type=int, default=256, help="frames per GPU per batch (configs[1])")
    ap.add_argument("--inner", type=int, default=64, help="batches per step (a step = inner x batch frames per GPU)")
    ap.add_argument("--store-rows", type=int, default=0, help="total store rows (default 10M at N=1, 100M at N>1)")
    ap.add_argument("--queries", type=int, default=10_000)
    ap.add_argument("--k", type=int, default=100)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the N=1 secondary lines (100M store, L/14, config 5, "
                                                             "scene-cut, library baseline)")
    ap.add_argument("--no-library-compile", action="store_true", help="skip torch.compile of the stock tower")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference(args)
        return

    import torch
    import torch.distributed as dist
    rank, local_rank, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this package has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import __graft_entry__ as ge
    pkg = ge.load_package()
    from oracle import search_oracle as so, vit_oracle as vo  # cpu_baseline leg only (rank 0, N=1)
    peaks = load_peaks()

    line = {"metric": METRIC, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": encode_workload(args), "parallelism": f"dp{world}",
                       "step": f"{args.inner} consecutive batches of {args.batch} frames per GPU "
                               f"({args.inner * args.batch} frames): K steps are seconds of sustained work",
                       "l2": "8 distinct resident uint8 batches (308 MB) rotate; weights+activations exceed L2"}}
    extras = rank == 0 and world == 1 and args.workload == "both" and not args.no_extras
    if args.workload in ("both", "encode"):
        r = bench_encode(pkg, dev, rank, world, args.steps, args.warmup, args.batch, args.inner, peaks)
        line.update(value=round(r["value"], 1), ms_per_step=round(r["ms_per_step"], 4), e2e=r["e2e"],
                    gpu_launches=r["gpu_launches"], clocks=r["clocks"], roofline=r["roofline"],
                    timed_region_s=round(r["timed_region_s"], 3), parity=r["parity"])
    if args.workload in ("both", "search"):
        total = args.store_rows or (10_000_000 if world == 1 else 100_000_000)
        s_steps = max(3, min(args.steps, 10))
        line["search"] = bench_search(pkg, dev, rank, world, s_steps, 3, total, args.queries, args.k, peaks)
        if "parity" in line:
            line["parity"]["search"] = line["search"]["parity"]
        if args.workload == "search":
            s = line["search"]
            line.update(metric=s["metric"], unit=s["unit"], value=s["value"], ms_per_step=s["ms_per_step"],
                        steps=s["steps"], scaling=s["scaling"], dtype="f16", e2e=s["e2e"],
                        gpu_launches=s["gpu_launches"], roofline=s.get("roofline"), config=s["config"],
                        parity=s["parity"])
    if extras:
        # the N = 1 point of the strong-scaling curve of configs[3]: the 100 M-row store (102.4 GB) on one B200
        line["search_100m"] = bench_search(pkg, dev, rank, world, 3, 3, 100_000_000, args.queries, args.k, peaks,
                                           hbm_regime=False, label="queries/s top-100 over 100M-frame store, one GPU")
        # BASELINE config 5's tower (ViT-L/14, 768-d) as a secondary line + config 5 end to end
        line["vit_l14"] = bench_encode_other(pkg, dev, "ViT-L/14", 128, 40, 5, peaks)
        line["config5"] = bench_config5(pkg, dev, peaks)
        line["scenecut"] = bench_scenecut(pkg, dev, peaks)
        line["ingest_jpeg"] = bench_ingest_jpeg(pkg, dev, peaks)
        line["library_baseline"] = bench_library(dev, args.batch, args.k, not args.no_library_compile)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        fps, threads = cpu_encode(vo, n_batches=8)
        qps = cpu_search(so, 1_000_000, 4, args.k)
        line["cpu_baseline"] = {
            "value": round(fps, 2), "unit": "frames/s", "cores": threads, "kind": "port",
            "sample": "256 frames (8 batches of 32, the reference's batch size) through oracle/vit_oracle.py on the "
                      "host cores; the reference's `clip` dependency is an absent submodule",
            "search": {"value": round(qps / 10.0, 3), "unit": "queries/s",
                       "sample": f"4 queries top-{args.k} over 1 M fp32 rows ({qps:.2f} q/s), scaled linearly to 10 M"}}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
