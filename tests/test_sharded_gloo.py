"""N>1 host logic on CPU: world_size-2 `gloo` run of the sharded search's partition + exchange step.

The per-rank scan and the final merge are CUDA kernels (tests/test_search_gpu.py covers them, including
shard+merge == unsharded bit-exactly on one GPU).  Here the two ranks stand in for them with the numpy oracle so
that the part that only exists at N>1 -- `shard_bounds`, padding of short shards, the [P, Q, k] all-gather layout
of `exchange_topk` -- runs for real over a process group.
"""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

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
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        search = importlib.import_module(PKG + ".search")
        from oracle import search_oracle as so
        rng = np.random.default_rng(5)  # every rank draws the same store and queries
        raw = rng.standard_normal((N, 64)).astype(np.float32)
        q = rng.standard_normal((Q, 64)).astype(np.float32)
        q /= np.linalg.norm(q, axis=-1, keepdims=True)
        emb = so.normalise_store(raw)
        lo, hi = search.shard_bounds(N, world, rank)
        kk = min(k, hi - lo)
        idx, sc = so.batched_topk(emb[lo:hi], q, kk)                  # stand-in for the local CUDA scan
        s = torch.from_numpy(np.ascontiguousarray(sc, dtype=np.float32))
        i = torch.from_numpy(np.ascontiguousarray(idx, dtype=np.int64) + lo)  # row_offset = lo
        all_s, all_i = search.exchange_topk(s, i, k, world)
        assert all_s.shape == (world, Q, k) and all_i.shape == (world, Q, k)
        assert torch.equal(all_s[rank, :, :kk], s) and torch.equal(all_i[rank, :, :kk], i)
        if kk < k:
            assert torch.all(all_s[rank, :, kk:] == float("-inf")) and torch.all(all_i[rank, :, kk:] == -1)
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), s=all_s.numpy(), i=all_i.numpy(), lo=lo, hi=hi)
    finally:
        dist.destroy_process_group()


def _run(tmp_path, N, Q, k, world=2):
    port = _free_port()
    mp.spawn(_worker, args=(world, port, N, Q, k, str(tmp_path)), nprocs=world, join=True)
    return [np.load(tmp_path / f"rank{r}.npz") for r in range(world)]


def _check(outs, N, Q, k):
    sys.path.insert(0, ROOT)
    from oracle import search_oracle as so
    rng = np.random.default_rng(5)
    raw = rng.standard_normal((N, 64)).astype(np.float32)
    q = rng.standard_normal((Q, 64)).astype(np.float32)
    q /= np.linalg.norm(q, axis=-1, keepdims=True)
    ref_i, ref_s = so.batched_topk(so.normalise_store(raw), q, min(k, N))
    # every rank holds the same gathered lists, shard r in slot r, and the shards tile [0, N)
    assert np.array_equal(outs[0]["s"], outs[1]["s"]) and np.array_equal(outs[0]["i"], outs[1]["i"])
    assert int(outs[0]["lo"]) == 0 and int(outs[0]["hi"]) == int(outs[1]["lo"]) and int(outs[1]["hi"]) == N
    s, i = outs[0]["s"], outs[0]["i"]
    for j in range(Q):  # what evr_topk_merge computes: best k of the P*k candidates, ties by ascending row
        fs, fi = s[:, j].ravel(), i[:, j].ravel()
        order = np.lexsort((fi, -fs))[:min(k, N)]
        assert so.topk_equal_up_to_ties(fi[order], fs[order], ref_i[j], ref_s[j], 1e-6)


def test_two_rank_exchange_matches_unsharded(tmp_path):
    N, Q, k = 1001, 5, 10  # odd N: shards of 501 and 500 rows
    _check(_run(tmp_path, N, Q, k), N, Q, k)


def test_two_rank_short_shards_are_padded(tmp_path):
    N, Q, k = 13, 3, 10  # shards of 7 and 6 rows < k: padded with (-inf, -1), merged list still exact
    _check(_run(tmp_path, N, Q, k), N, Q, k)


def test_shard_bounds_tile_the_store():
    import importlib
    sys.path.insert(0, ROOT)
    search = importlib.import_module(PKG + ".search")
    for n in (0, 1, 7, 1000, 100_000_000):
        for world in (1, 2, 4, 8):
            spans = [search.shard_bounds(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
