"""Concatenated multi-video store: global row -> (video, frame) identity and search across videos,
against the numpy oracle over the concatenated, per-video-normalised stores; row-sharded == unsharded."""
import importlib
import json

import numpy as np
import pytest
import torch

from conftest import PKG
from oracle import search_oracle as so

pytestmark = pytest.mark.gpu


def _site(tmp_path, rng):
    specs = [("clip_b", 700, np.float32, 3.0), ("clip_a", 1300, np.float16, 9.0), ("short", 40, np.float32, 1.0)]
    mapping, raws, names = {}, [], []
    for video, n, dt, scale in specs:
        raw = (rng.standard_normal((n, 512)) * scale).astype(dt)
        np.save(tmp_path / f"{video}_embeddings.npy", raw)
        frames = [{"filepath": f"/frames/{video}/{i * 7}.jpg", "frameidx": i * 7} for i in range(n)]
        if video == "short":
            frames = frames[:-3]                      # metadata shorter than the store: rows beyond it are dropped
        json.dump(frames, open(tmp_path / f"{video}_metadata.json", "w"))
        mapping[video] = {"metadata_file": str(tmp_path / f"{video}_metadata.json"),
                          "embeddings_file": str(tmp_path / f"{video}_embeddings.npy"), "video_path": "", "frames_dir": ""}
        keep = len(frames)
        raws.append(so.normalise_store(raw[:keep]).astype(np.float32))
        names += [(video, f"{i * 7}.jpg") for i in range(keep)]
    mapping["ghost"] = {"metadata_file": str(tmp_path / "ghost.json"), "embeddings_file": str(tmp_path / "ghost.npy")}
    return mapping, np.concatenate(raws), names


def test_search_across_videos_matches_oracle(pkg, cuda, tmp_path):
    rng = np.random.default_rng(12)
    mapping, emb, names = _site(tmp_path, rng)
    store = pkg.store.MultiVideoStore.from_mapping(mapping, device=cuda, chunk_rows=512)
    assert store.N == len(names) == 700 + 1300 + 37 and store.videos == ["clip_b", "clip_a", "short"]
    assert store.offsets == [0, 700, 2000, 2037] and store.video_rows("clip_a") == (700, 2000)
    for g in (0, 699, 700, 1999, 2000, 2036):
        video, frame, local = store.locate(g)
        assert (video, frame) == names[g] and local == g - store.offsets[store.videos.index(video)]
    with pytest.raises(IndexError):
        store.locate(2037)
    q = rng.standard_normal((5, 512)).astype(np.float32)
    q /= np.linalg.norm(q, axis=-1, keepdims=True)
    sc, idx = store.search(torch.from_numpy(q), 50)
    ref_i, ref_s = so.batched_topk(emb, q, 50)
    for j in range(5):
        assert so.topk_equal_up_to_ties(idx[j].cpu().numpy(), sc[j].cpu().numpy(), ref_i[j], ref_s[j], 2.0 ** -10)
    hits = store.search_frames(q[0], 10)
    assert [(v, f) for v, f, _ in hits] == [names[i] for i in idx[0, :10].cpu().numpy()]
    assert all(isinstance(s, float) for _, _, s in hits) and hits[0][2] >= hits[-1][2]


def test_row_sharded_store_equals_whole(pkg, cuda, tmp_path):
    rng = np.random.default_rng(13)
    mapping, emb, names = _site(tmp_path, rng)
    whole = pkg.store.MultiVideoStore.from_mapping(mapping, device=cuda)
    q = torch.from_numpy(rng.standard_normal((7, 512)).astype(np.float32))
    ws, wi = whole.search(q, 30)
    parts_s, parts_i, covered = [], [], 0
    for r in range(3):                                 # three ranks emulated on one GPU; shards cut through videos
        shard = pkg.store.MultiVideoStore.from_mapping(mapping, device=cuda, rank=r, world=3, chunk_rows=300)
        assert shard.lo == covered and shard.local.N == shard.hi - shard.lo and shard.local.row_offset == shard.lo
        covered = shard.hi
        assert torch.equal(shard.local.rows, whole.local.rows[shard.lo:shard.hi])   # same normalised rows, bit for bit
        s, i = shard.local.search(q, 30)
        parts_s.append(s)
        parts_i.append(i)
    assert covered == whole.N
    ms, mi = pkg.search.merge_topk(torch.stack(parts_s), torch.stack(parts_i))
    assert torch.equal(ms, ws) and torch.equal(mi, wi)


def test_empty_mapping_is_an_error(pkg, cuda, tmp_path):
    with pytest.raises(ValueError):
        pkg.store.MultiVideoStore.from_mapping({"ghost": {"metadata_file": "x", "embeddings_file": str(tmp_path / "no.npy")}},
                                               device=cuda)
