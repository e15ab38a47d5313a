"""Data-parallel ingest on CPU: world_size-2 (and 3) `gloo` runs of `EmbeddingService.extract_and_save_embeddings_from_folder`.

What only exists at N > 1 -- contiguous rank slices of the `sorted()` frame list (embedding_service.py:452), the pre-sized
shared `.npy`, the two barriers, the rank-0 metadata stamp -- runs for real over a process group.  The tower itself is a
CUDA kernel, so `_encode_frames` is replaced by a deterministic per-file stand-in (a row seeded by the file name); the
GPU twin with the real tower is tests/test_services_gpu.py::test_sharded_ingest_equals_single (one GPU, two slices) and
tests/test_sharded_nccl.py (one process per GPU).  Done = the file equals the single-process file byte for byte.
"""
import json
import os
import socket
import sys

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = "event-retrival-in-video-learning-transferable-visual-model-from-supervised-natural-language_b200"
NAMES = [f"{i}.jpg" for i in (10074, 1061, 7, 300, 25, 42, 9000, 11, 512, 77, 3)] + ["skip.JPG", "notes.txt"]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _fake_service(services, tmp, dtype):
    class Stub(services.EmbeddingService):
        def _encode_frames(self, folder_path, frame_files, key):
            rows = [np.random.default_rng(int(os.path.splitext(f)[0])).standard_normal(16) * 3 for f in frame_files]
            bad = [i for i, f in enumerate(frame_files) if f == "42.jpg"]   # one undecodable frame -> zero row
            return np.asarray(rows, dtype=dtype).reshape(len(frame_files), 16), bad

    svc = Stub.__new__(Stub)
    ps = services.PathService(os.path.join(tmp, "frames"), os.path.join(tmp, "metadata"), os.path.join(tmp, "embedding"))
    svc.path_service, svc.active_model, svc.finetuned_model = ps, "original", None
    return svc


def _layout(tmp):
    for d in ("frames/vid", "metadata", "embedding"):
        os.makedirs(os.path.join(tmp, d), exist_ok=True)
    for n in NAMES:
        open(os.path.join(tmp, "frames/vid", n), "w").write("x")
    meta = os.path.join(tmp, "metadata", "vid_metadata.json")
    json.dump([{"filepath": f"/f/{n}"} for n in sorted(n for n in NAMES if n.endswith(".jpg"))], open(meta, "w"))
    mapping = {"vid": {"metadata_file": meta, "embeddings_file": os.path.join(tmp, "embedding", "vid_embeddings.npy"),
                       "video_path": "", "frames_dir": os.path.join(tmp, "frames/vid")}}
    json.dump(mapping, open(os.path.join(tmp, "metadata", "video_mapping.json"), "w"))


def _worker(rank, world, port, tmp, dtype):
    import importlib
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        services = importlib.import_module(PKG + ".services")
        svc = _fake_service(services, tmp, np.dtype(dtype))
        path = svc.extract_and_save_embeddings_from_folder(os.path.join(tmp, "frames/vid"), None, "vid")
        assert path == os.path.join(tmp, "embedding", "vid_embeddings.npy")
    finally:
        dist.destroy_process_group()


def _single(tmp, dtype):
    import importlib
    sys.path.insert(0, ROOT)
    services = importlib.import_module(PKG + ".services")
    svc = _fake_service(services, tmp, np.dtype(dtype))
    return svc.extract_and_save_embeddings_from_folder(os.path.join(tmp, "frames/vid"), None, "vid", rank=0, world=1)


def _run(tmp_path, world, dtype):
    one, many = str(tmp_path / "one"), str(tmp_path / "many")
    _layout(one)
    _layout(many)
    single = _single(one, dtype)
    mp.spawn(_worker, args=(world, _free_port(), many, dtype), nprocs=world, join=True)
    a = open(single, "rb").read()
    b = open(os.path.join(many, "embedding", "vid_embeddings.npy"), "rb").read()
    assert a == b, "the rank-sliced file differs from the single-process file"
    rows = np.load(single)
    order = sorted(n for n in NAMES if n.endswith(".jpg"))          # lexicographic, case-sensitive filter
    assert rows.shape == (len(order), 16) and order[:2] == ["10074.jpg", "1061.jpg"]
    assert np.all(rows[order.index("42.jpg")] == 0)
    norms = np.linalg.norm(np.delete(rows, order.index("42.jpg"), 0).astype(np.float32), axis=-1)
    assert np.allclose(norms, 1.0, atol=2e-3)
    for r in range(world):                                          # rank 0 (only) stamped the metadata, once
        meta = json.load(open(os.path.join(many, "metadata", "vid_metadata.json")))
        assert all(item["embedding_model"] == "original" for item in meta)


def test_two_ranks_write_the_single_process_file(tmp_path):
    _run(tmp_path, 2, "float32")


def test_three_ranks_fp16_rows(tmp_path):
    _run(tmp_path, 3, "float16")     # 11 frames over 3 ranks: slices of 4, 4, 3


def test_more_ranks_than_frames_is_fine(tmp_path):
    import importlib
    sys.path.insert(0, ROOT)
    services = importlib.import_module(PKG + ".services")
    path = str(tmp_path / "x.npy")
    rows = np.arange(6, dtype=np.float32).reshape(2, 3)
    # emulate 3 ranks in one process: rank 2 holds no row
    services.shard_rows_to_npy(path, 2, 0, rows[:1], 0, 3)
    services.shard_rows_to_npy(path + ".tmp.npy", 2, 0, rows[:1], 0, 1)
    mm = np.load(path, mmap_mode="r+")
    mm[1:2] = rows[1:]
    mm.flush()
    assert np.array_equal(np.load(path), rows)
    import pytest
    with pytest.raises(ValueError, match="process group"):
        services._dist_context(1, 2, None)
