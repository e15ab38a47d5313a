"""Row-sharded search on real GPUs: one process per GPU, NCCL all-gather of the per-rank lists, on-device merge
(ShardedFrameStore).  Needs >= 2 CUDA devices; on a one-GPU box the test is skipped (the one-GPU emulation in
test_search_gpu.py and the gloo test cover the same logic)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = "event-retrival-in-video-learning-transferable-visual-model-from-supervised-natural-language_b200"


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, N, Q, k, out_dir):
    import importlib
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        search = importlib.import_module(PKG + ".search")
        rng = np.random.default_rng(17)          # every rank draws the same store and queries
        raw = rng.standard_normal((N, 512)).astype(np.float32)
        q = rng.standard_normal((Q, 512)).astype(np.float32)
        q /= np.linalg.norm(q, axis=-1, keepdims=True)
        lo, hi = search.shard_bounds(N, world, rank)
        local = search.FrameStore.from_numpy(raw[lo:hi], device=dev, row_offset=lo)
        sc, idx = search.ShardedFrameStore(local).search(torch.from_numpy(q).to(dev), k)
        torch.cuda.synchronize()
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), s=sc.cpu().numpy(), i=idx.cpu().numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("N,Q,k", [(50_001, 9, 100), (150, 3, 100)])
def test_nccl_sharded_search_matches_oracle(tmp_path, N, Q, k):
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs at least 2 CUDA devices")
    sys.path.insert(0, ROOT)
    from oracle import search_oracle as so
    mp.spawn(_worker, args=(world, _free_port(), N, Q, k, str(tmp_path)), nprocs=world, join=True)
    outs = [np.load(tmp_path / f"rank{r}.npz") for r in range(world)]
    for o in outs[1:]:                                   # every rank holds the same merged lists
        assert np.array_equal(o["s"], outs[0]["s"]) and np.array_equal(o["i"], outs[0]["i"])
    rng = np.random.default_rng(17)
    raw = rng.standard_normal((N, 512)).astype(np.float32)
    q = rng.standard_normal((Q, 512)).astype(np.float32)
    q /= np.linalg.norm(q, axis=-1, keepdims=True)
    kk = min(k, N)
    ref_i, ref_s = so.batched_topk(so.normalise_store(raw), q, kk)
    s, i = outs[0]["s"], outs[0]["i"]
    assert s.shape == (Q, k)
    for j in range(Q):
        assert so.topk_equal_up_to_ties(i[j, :kk], s[j, :kk], ref_i[j], ref_s[j], 2.0 ** -10)
        assert (i[j, kk:] == -1).all() and np.isneginf(s[j, kk:]).all()   # fewer rows than k: padded tail


def test_one_process_two_devices(pkg):
    """A single process that touches two GPUs: per-device kernel attributes / SM counts must not leak across devices."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 CUDA devices")
    sys.path.insert(0, ROOT)
    from oracle import search_oracle as so
    rng = np.random.default_rng(3)
    raw = rng.standard_normal((20_000, 512)).astype(np.float32)
    q = rng.standard_normal((150, 512)).astype(np.float32)
    q /= np.linalg.norm(q, axis=-1, keepdims=True)
    ref_i, ref_s = so.batched_topk(so.normalise_store(raw), q, 200)
    sd = pkg.clip.random_visual_state_dict("ViT-B/32", seed=1)
    frames = torch.randint(0, 256, (5, 224, 224, 3), dtype=torch.uint8)
    embs = []
    for dev in (torch.device("cuda:0"), torch.device("cuda:1")):
        store = pkg.search.FrameStore.from_numpy(raw, device=dev)
        sc, idx = store.search(torch.from_numpy(q).to(dev), 200)     # k > 128: the large select configuration
        for j in (0, 77, 149):
            assert so.topk_equal_up_to_ties(idx[j].cpu().numpy(), sc[j].cpu().numpy(), ref_i[j], ref_s[j], 2.0 ** -10)
        enc = pkg.vit.VisionEncoder.from_state_dict(sd, "ViT-B/32", device=dev, max_batch=8, prefix="visual.")
        embs.append(enc.encode(frames.to(dev), normalise=True).cpu())
    assert torch.equal(embs[0], embs[1])
