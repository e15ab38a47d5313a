"""`evr_topk_search_static`: the fused scan + top-k with its scratch in a static device array (no allocation in the call).

The entry point has no counterpart in the reference (added for callers that cannot
call cudaMalloc).  It must return exactly what `evr_topk_search` returns with a caller-owned workspace -- same kernels,
same tie rule, so indices AND scores are compared bit for bit (the tolerance of tests/test_search_gpu.py for identical
operands) -- on two calls in a row (the array keeps the previous call's contents), must allocate nothing through torch
while it runs, and must refuse a request whose scratch exceeds EVR_TOPK_STATIC_SCRATCH_BYTES with a clear error.

Collected last on purpose: nothing here depends on the tests before it.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

STATIC_BYTES = 80 << 20     # EVR_TOPK_STATIC_SCRATCH_BYTES, include/evr.h


def _unit_rows(n, d, seed, device):
    g = torch.Generator(device=device).manual_seed(seed)
    x = torch.randn((n, d), generator=g, device=device)
    return (x / x.norm(dim=-1, keepdim=True)).half().contiguous()


def _search(pkg, store, queries, k, row_offset):
    """existing entry point, caller-owned workspace (torch allocates it, as FrameStore does)"""
    L = pkg._lib
    N, D = store.shape
    Q = queries.shape[0]
    need = ctypes.c_size_t()
    L.check(L.lib.evr_topk_workspace_bytes(N, D, Q, k, ctypes.byref(need)), "evr_topk_workspace_bytes")
    ws = torch.empty(need.value, dtype=torch.uint8, device=store.device)
    scores = torch.full((Q, k), float("nan"), device=store.device)
    idx = torch.full((Q, k), -7, dtype=torch.int64, device=store.device)
    L.check(L.lib.evr_topk_search(L.ptr(store), N, D, row_offset, L.ptr(queries), Q, k, L.ptr(scores), L.ptr(idx),
                                  L.ptr(ws), need.value, L.current_stream_ptr()), "evr_topk_search")
    torch.cuda.synchronize()
    return scores, idx, need.value


def _search_static(pkg, store, queries, k, row_offset, scores, idx):
    L = pkg._lib
    N, D = store.shape
    rc = L.lib.evr_topk_search_static(L.ptr(store), N, D, row_offset, L.ptr(queries), queries.shape[0], k, L.ptr(scores),
                                      L.ptr(idx), L.current_stream_ptr())
    L.check(rc, "evr_topk_search_static")
    torch.cuda.synchronize()


def test_static_scratch_scan_equals_the_workspace_scan_twice_in_a_row(pkg, cuda):
    store = _unit_rows(300_000, 512, 41, cuda)      # large enough for the sample pass and several slices per SM
    small = _unit_rows(5_000, 512, 42, cuda)        # no sample pass, a different plan in the same array afterwards
    calls = [(store, _unit_rows(16, 512, 43, cuda), 50, 0),
             (store, _unit_rows(128, 512, 44, cuda), 100, 1_000_000),     # the array holds call 1's lists, counts, flags
             (small, _unit_rows(3, 512, 45, cuda), 10, 0),
             (store, _unit_rows(16, 512, 43, cuda), 50, 0)]               # call 1 again
    for n, (st, q, k, off) in enumerate(calls):
        ref_scores, ref_idx, need = _search(pkg, st, q, k, off)
        assert need <= STATIC_BYTES
        scores = torch.full_like(ref_scores, float("nan"))
        idx = torch.full_like(ref_idx, -7)
        torch.cuda.synchronize()
        before = (torch.cuda.memory_allocated(cuda), torch.cuda.memory_reserved(cuda))
        _search_static(pkg, st, q, k, off, scores, idx)
        assert (torch.cuda.memory_allocated(cuda), torch.cuda.memory_reserved(cuda)) == before   # nothing through torch
        assert torch.equal(idx, ref_idx), f"call {n}: indices differ"
        assert torch.equal(scores, ref_scores), f"call {n}: scores differ"
        # and against plain torch on the same operands: the top-1 of every query
        best = (q.float() @ st.float().t()).max(dim=1).values
        assert torch.all((scores[:, 0] - best).abs() <= 2.0 ** -10)       # the epsilon of tests/test_search_gpu.py
        assert torch.all(idx >= off) and torch.all(idx < off + st.shape[0])


def test_static_scratch_scan_rejects_what_does_not_fit(pkg, cuda):
    store = _unit_rows(70_000, 512, 46, cuda)
    q = _unit_rows(1024, 512, 47, cuda)             # 8 query tiles: ~196 MiB of candidate lists
    scores = torch.empty((1024, 100), device=cuda)
    idx = torch.empty((1024, 100), dtype=torch.int64, device=cuda)
    with pytest.raises(pkg._lib.EvrError, match="bytes of scratch, the static array holds"):
        _search_static(pkg, store, q, 100, 0, scores, idx)
    with pytest.raises(pkg._lib.EvrError, match="null pointer"):
        L = pkg._lib
        L.check(L.lib.evr_topk_search_static(None, 10, 512, 0, L.ptr(q), 1, 1, L.ptr(scores), L.ptr(idx),
                                             L.current_stream_ptr()), "evr_topk_search_static")
    # the failed calls leave the entry point usable
    ref_scores, ref_idx, _ = _search(pkg, store, q[:8], 20, 0)
    s2, i2 = torch.empty_like(ref_scores), torch.empty_like(ref_idx)
    _search_static(pkg, store, q[:8], 20, 0, s2, i2)
    assert torch.equal(i2, ref_idx) and torch.equal(s2, ref_scores)


def test_static_scratch_scan_under_graph_capture(pkg, cuda):
    """The use the entry point was asked for: captured into a CUDA graph (global capture mode: a cudaMalloc anywhere in
    the call would fail the capture) and replayed with new query contents in the same buffers.  Every replay runs the
    captured clears of the counters and flags first, so it must equal a fresh `evr_topk_search` each time."""
    store = _unit_rows(120_000, 512, 51, cuda)
    q = _unit_rows(32, 512, 52, cuda)
    k = 25
    scores = torch.empty((32, k), device=cuda)
    idx = torch.empty((32, k), dtype=torch.int64, device=cuda)
    _search_static(pkg, store, q, k, 7, scores, idx)            # set-up outside the capture (symbol address, attributes)
    L = pkg._lib
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        L.check(L.lib.evr_topk_search_static(L.ptr(store), store.shape[0], 512, 7, L.ptr(q), 32, k, L.ptr(scores),
                                             L.ptr(idx), L.current_stream_ptr()), "evr_topk_search_static")
    for seed in (53, 54, 55):
        q.copy_(_unit_rows(32, 512, seed, cuda))
        scores.fill_(float("nan"))
        idx.fill_(-7)
        graph.replay()
        torch.cuda.synchronize()
        ref_scores, ref_idx, _ = _search(pkg, store, q, k, 7)
        assert torch.equal(idx, ref_idx) and torch.equal(scores, ref_scores), seed
