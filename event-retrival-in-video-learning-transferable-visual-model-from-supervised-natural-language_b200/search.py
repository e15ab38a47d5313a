"""HBM-resident frame-embedding store + fused cosine scan / top-k.

Replaces the numpy scan of Backend/services/embedding_service.py:
  get_embeddings :186-217 (load + L2 normalise)          -> FrameStore.from_numpy / from_npy
  search_top_frames :314-320, by_image :365-371          -> FrameStore.search  (evr_topk_search)
  extract_query_confidence :274-279                      -> FrameStore.rescore (evr_rescore)
Row-sharded multi-GPU search: every rank scans its own rows, the per-rank (score, idx)[Q, k]
lists are all-gathered over NCCL and merged on-device (evr_topk_merge).
"""
import ctypes

import numpy as np
import torch

from . import _lib
from ._lib import HandleGuard, lib, check, ptr, current_stream_ptr

MAX_FUSED_K = 1024


def _as_query_tensor(queries, device, D):
    if isinstance(queries, np.ndarray):
        queries = torch.from_numpy(np.ascontiguousarray(queries))
    q = queries.to(device=device)
    if q.dim() == 1:
        q = q.unsqueeze(0)
    if q.dim() != 2 or q.shape[1] != D:
        raise ValueError(f"queries must be [Q, {D}], got {tuple(q.shape)}")
    return q.to(torch.float16).contiguous()


class FrameStore:
    """[N, D] unit-norm fp16 rows resident in HBM (the layout the scan kernel streams with TMA)."""

    def __init__(self, rows_fp16, row_offset=0):
        assert rows_fp16.is_cuda and rows_fp16.dtype == torch.float16 and rows_fp16.is_contiguous()
        self.rows = rows_fp16
        self.N, self.D = rows_fp16.shape
        self.row_offset = int(row_offset)
        self.device = rows_fp16.device
        self._ws = None
        self._guard = HandleGuard()  # the candidate workspace is shared by every search on this store

    # ---- construction -------------------------------------------------------------------------
    @classmethod
    def from_tensor(cls, raw, device=None, row_offset=0):
        """raw: [N, D] fp32 or fp16 tensor (normalised or not) -> normalised fp16 store on `device`."""
        device = torch.device(device if device is not None else (raw.device if raw.is_cuda else "cuda"))
        if raw.dim() != 2:
            raise ValueError("store must be 2-D [N, D]")
        if raw.dtype not in (torch.float32, torch.float16):
            raw = raw.to(torch.float32)
        src = raw.to(device).contiguous()
        N, D = src.shape
        dst = torch.empty((N, D), dtype=torch.float16, device=device)
        if N > 0:
            with torch.cuda.device(device):
                check(lib.evr_store_prepare(ptr(src), 0 if src.dtype == torch.float32 else 1, N, D, ptr(dst),
                                            current_stream_ptr()), "evr_store_prepare")
        return cls(dst, row_offset)

    @classmethod
    def from_numpy(cls, raw, device="cuda", row_offset=0):
        raw = np.ascontiguousarray(raw)
        if raw.dtype not in (np.float32, np.float16):
            raw = raw.astype(np.float32)
        return cls.from_tensor(torch.from_numpy(raw), device, row_offset)

    @classmethod
    def from_npy(cls, path, device="cuda", row_offset=0):
        """Loads a `<video>_embeddings.npy` file ([N, D] <f2 / <f4, embedding_service.py:505 layout)."""
        return cls.from_numpy(np.load(path), device, row_offset)

    # ---- scan -----------------------------------------------------------------------------------
    def _workspace(self, Q, k):
        need = ctypes.c_size_t(0)
        check(lib.evr_topk_workspace_bytes(self.N, self.D, Q, k, ctypes.byref(need)), "evr_topk_workspace_bytes")
        if self._ws is None or self._ws.numel() < need.value:
            self._ws = torch.empty(need.value, dtype=torch.uint8, device=self.device)
        return self._ws

    def search(self, queries, k):
        """queries [Q, D] (any float dtype) -> (scores fp32 [Q, k'] desc, idx int64 [Q, k']), k' = min(k, N).

        idx are global row ids (row_offset + local row)."""
        q = _as_query_tensor(queries, self.device, self.D)
        Q = q.shape[0]
        k_eff = min(int(k), self.N)
        if Q == 0 or k_eff <= 0:
            kk = max(k_eff, 0)
            return (torch.empty((Q, kk), dtype=torch.float32, device=self.device),
                    torch.empty((Q, kk), dtype=torch.int64, device=self.device))
        if k_eff > MAX_FUSED_K:
            return self._search_large_k(q, k_eff)
        out_s = torch.empty((Q, k_eff), dtype=torch.float32, device=self.device)
        out_i = torch.empty((Q, k_eff), dtype=torch.int64, device=self.device)
        with torch.cuda.device(self.device), self._guard:
            ws = self._workspace(Q, k_eff)
            check(lib.evr_topk_search(ptr(self.rows), self.N, self.D, self.row_offset, ptr(q), Q, k_eff, ptr(out_s),
                                      ptr(out_i), ptr(ws), ws.numel(), current_stream_ptr()), "evr_topk_search")
        return out_s, out_i

    def _search_large_k(self, q, k):
        """k > 1024 (the reference takes any top_k: it sorts all N scores, embedding_service.py:317-320): a score pass
        over the store per chunk of 8 queries + a per-query radix select / sort on the device
        (evr_topk_search_large).  Same outputs and tie rule as the fused path."""
        Q = q.shape[0]
        out_s = torch.empty((Q, k), dtype=torch.float32, device=self.device)
        out_i = torch.empty((Q, k), dtype=torch.int64, device=self.device)
        need = ctypes.c_size_t(0)
        check(lib.evr_topk_large_workspace_bytes(self.N, Q, k, ctypes.byref(need)), "evr_topk_large_workspace_bytes")
        with torch.cuda.device(self.device), self._guard:
            if self._ws is None or self._ws.numel() < need.value:
                self._ws = torch.empty(need.value, dtype=torch.uint8, device=self.device)
            check(lib.evr_topk_search_large(ptr(self.rows), self.N, self.D, self.row_offset, ptr(q), Q, k, ptr(out_s),
                                            ptr(out_i), ptr(self._ws), self._ws.numel(), current_stream_ptr()),
                  "evr_topk_search_large")
        return out_s, out_i

    def rescore(self, query, rows):
        """cosine of `query` [D] with the given local rows -> fp32 [R] (embedding_service.py:274-279)."""
        q = _as_query_tensor(query, self.device, self.D)[0].contiguous()
        r = torch.as_tensor(rows, dtype=torch.int64, device=self.device).contiguous()
        out = torch.empty(r.numel(), dtype=torch.float32, device=self.device)
        if r.numel():
            with torch.cuda.device(self.device):
                check(lib.evr_rescore(ptr(self.rows), self.N, self.D, ptr(q), ptr(r), r.numel(), ptr(out),
                                      current_stream_ptr()), "evr_rescore")
        return out


def merge_topk(scores, idx):
    """scores [P, Q, k] fp32, idx [P, Q, k] int64 (each list sorted or not) -> merged ([Q, k], [Q, k])."""
    assert scores.is_cuda and scores.dtype == torch.float32 and idx.dtype == torch.int64
    scores, idx = scores.contiguous(), idx.contiguous()
    P, Q, k = scores.shape
    out_s = torch.empty((Q, k), dtype=torch.float32, device=scores.device)
    out_i = torch.empty((Q, k), dtype=torch.int64, device=scores.device)
    with torch.cuda.device(scores.device):
        if k <= MAX_FUSED_K:
            check(lib.evr_topk_merge(ptr(scores), ptr(idx), P, Q, k, ptr(out_s), ptr(out_i), current_stream_ptr()),
                  "evr_topk_merge")
        else:
            kp = 1 << (k - 1).bit_length()
            scratch = torch.empty(Q * kp * 8, dtype=torch.uint8, device=scores.device)
            check(lib.evr_topk_merge_large(ptr(scores), ptr(idx), P, Q, k, ptr(out_s), ptr(out_i), ptr(scratch),
                                           scratch.numel(), current_stream_ptr()), "evr_topk_merge_large")
    return out_s, out_i


def shard_bounds(n_rows, world_size, rank):
    """Contiguous row partition of the store: rank r owns [lo, hi)."""
    base, rem = divmod(n_rows, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class ShardedFrameStore:
    """Row-partitioned store: one FrameStore per rank + all-gather + on-device k-way merge."""

    def __init__(self, local_store, group=None):
        import torch.distributed as dist
        self.local = local_store
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1

    def search(self, queries, k):
        s, i = self.local.search(queries, k)
        if self.world == 1:
            return s, i
        all_s, all_i = exchange_topk(s, i, k, self.world, self.group)
        return merge_topk(all_s, all_i)


def exchange_topk(s, i, k, world, group=None):
    """The one exchange step of the sharded search: every rank's (scores [Q, k'], idx [Q, k']) list, padded to k
    with (-inf, -1) when its shard holds fewer than k rows, is all-gathered into [P, Q, k] on every rank
    (NCCL over NVLink on the GPUs; the same code runs under gloo in tests/test_sharded_gloo.py)."""
    import torch.distributed as dist
    Q, kk = s.shape
    if kk < k:
        s = torch.cat([s, torch.full((Q, k - kk), float("-inf"), dtype=s.dtype, device=s.device)], 1)
        i = torch.cat([i, torch.full((Q, k - kk), -1, dtype=torch.int64, device=i.device)], 1)
    all_s = torch.empty((world, Q, k), dtype=torch.float32, device=s.device)
    all_i = torch.empty((world, Q, k), dtype=torch.int64, device=s.device)
    # output viewed as the dim-0 concatenation [P*Q, k] of the rank inputs (the form every backend accepts)
    dist.all_gather_into_tensor(all_s.view(world * Q, k), s.contiguous(), group=group)
    dist.all_gather_into_tensor(all_i.view(world * Q, k), i.contiguous(), group=group)
    return all_s, all_i
