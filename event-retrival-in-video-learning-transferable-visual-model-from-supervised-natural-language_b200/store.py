"""Concatenated multi-video frame store with a global row -> (video, frame) table.

The reference searches one video's `.npy` per request (`video_name=None` -> the first mapping entry,
path_service.py:47-93), re-reads the metadata JSON on every request (query_strategies.py:59-63) and walks it with
O(N) `list.index` / `next(...)` per candidate (embedding_service.py:264, query_strategies.py:95).  The 10 M / 100 M-frame
stores of BASELINE.json are therefore a new object: every video of `video_mapping.json` laid end to end in one
HBM-resident fp16 store (normalised on the device, evr_store_prepare), plus
  offsets[v]            first global row of video v (mapping order)
  names                 frame file name of every row (metadata order = .npy row order, data_service.py:48)
so a hit (global row) maps to (video, frame name) in O(log V) with no JSON access.  Row-sharded use: rank r keeps only
global rows [lo, hi) (search.shard_bounds); files are memory-mapped, so a rank reads just its slice.
"""
import bisect
import json
import os

import numpy as np
import torch

from ._lib import lib, check, ptr, current_stream_ptr
from .search import FrameStore, ShardedFrameStore, shard_bounds


def _frame_names(metadata_path):
    with open(metadata_path, 'r', encoding='utf-8') as f:
        samples = json.load(f)
    return [os.path.basename(s["filepath"]) for s in samples if "filepath" in s]


class MultiVideoStore:
    def __init__(self, local_store, videos, offsets, names, lo, hi, group=None):
        self.local = local_store                      # FrameStore over global rows [lo, hi)
        self.videos, self.offsets, self.names = videos, offsets, names
        self.lo, self.hi = lo, hi
        self.N = offsets[-1]
        self._sharded = ShardedFrameStore(local_store, group)

    @classmethod
    def from_mapping(cls, video_data_mapping, device="cuda", rank=0, world=1, group=None, chunk_rows=1 << 20):
        """video_data_mapping: the dict of video_mapping.json ({video: {"metadata_file", "embeddings_file", ...}})."""
        videos, arrays, names, offsets = [], [], [], [0]
        D = None
        for video, entry in video_data_mapping.items():
            path = os.path.normpath(entry["embeddings_file"])
            if not os.path.exists(path):
                print(f"Warning: Embeddings file not found: {path}")
                continue
            arr = np.load(path, mmap_mode="r")
            if arr.ndim != 2 or arr.shape[0] == 0:
                continue
            if D is None:
                D = arr.shape[1]
            if arr.shape[1] != D:
                raise ValueError(f"{video}: embedding width {arr.shape[1]} differs from {D} (stores of different models)")
            frames = _frame_names(os.path.normpath(entry["metadata_file"]))
            n = min(len(frames), arr.shape[0])
            if len(frames) != arr.shape[0]:   # row <-> frame identity silently depends on equal lengths
                print(f"Warning: {video}: {arr.shape[0]} embedding rows but {len(frames)} metadata frames; using {n}")
            videos.append(video)
            arrays.append(arr[:n])
            names.extend(frames[:n])
            offsets.append(offsets[-1] + n)
        if not videos:
            raise ValueError("no usable video in the mapping")
        N = offsets[-1]
        lo, hi = shard_bounds(N, world, rank)
        device = torch.device(device)
        rows = torch.empty((hi - lo, D), dtype=torch.float16, device=device)
        with torch.cuda.device(device):
            for v, arr in enumerate(arrays):          # normalise every overlapping piece straight into its slice
                a, b = max(lo, offsets[v]), min(hi, offsets[v + 1])
                for c0 in range(a, b, chunk_rows):
                    c1 = min(b, c0 + chunk_rows)
                    piece = np.array(arr[c0 - offsets[v]:c1 - offsets[v]])   # copy out of the read-only memmap
                    if piece.dtype not in (np.float16, np.float32):
                        piece = piece.astype(np.float32)
                    src = torch.from_numpy(piece).to(device)
                    check(lib.evr_store_prepare(ptr(src), 0 if src.dtype == torch.float32 else 1, c1 - c0, D,
                                                ptr(rows[c0 - lo:c1 - lo]), current_stream_ptr()), "evr_store_prepare")
            torch.cuda.synchronize(device)
        return cls(FrameStore(rows, row_offset=lo), videos, offsets, np.asarray(names, dtype=object), lo, hi, group)

    @classmethod
    def from_path_service(cls, path_service, **kw):
        return cls.from_mapping(path_service.video_data_mapping, **kw)

    # ---- lookups ------------------------------------------------------------------------------------
    def locate(self, global_row):
        """global row -> (video, frame file name, row inside the video)."""
        g = int(global_row)
        if not 0 <= g < self.N:
            raise IndexError(g)
        v = bisect.bisect_right(self.offsets, g) - 1
        return self.videos[v], self.names[g], g - self.offsets[v]

    def video_rows(self, video):
        v = self.videos.index(video)
        return self.offsets[v], self.offsets[v + 1]

    # ---- search -------------------------------------------------------------------------------------
    def search(self, queries, k):
        """-> (scores [Q, k'], global rows [Q, k']) over every video (all-gather + merge when sharded)."""
        return self._sharded.search(queries, k)

    def search_frames(self, query, k):
        """One query vector -> [(video, frame name, score)] best first, across all videos."""
        s, i = self.search(torch.as_tensor(np.asarray(query, dtype=np.float32).reshape(1, -1)), k)
        s, i = s[0].cpu().numpy(), i[0].cpu().numpy()
        out = []
        for score, row in zip(s, i):
            if row < 0:
                continue
            video, name, _ = self.locate(row)
            out.append((video, name, float(score)))
        return out
