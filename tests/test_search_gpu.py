"""GPU parity of the fused cosine scan + top-k (through the C ABI) against the numpy oracle, the golden fixtures made
from the reference's stored stores, and size-independent properties at BASELINE's full store size.

Tolerance: the store is held as unit-norm fp16 and queries are rounded to fp16, products are
accumulated in fp32 on the tensor cores; scores must agree with the fp32 numpy path within EPS = 2**-10 and the ranked
indices must be identical except for permutations among scores closer than EPS."""
import os

import numpy as np
import pytest
import torch

from oracle import search_oracle as so

pytestmark = pytest.mark.gpu
EPS = 2.0 ** -10
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "search_golden.npz")
STORES = ["video_test_3", "video_test_4", "Violence", "Violence_2"]


def _unit(x):
    return x / np.linalg.norm(x, axis=-1, keepdims=True)


def _check(pkg, cuda, raw, q, k, row_offset=0):
    store = pkg.search.FrameStore.from_numpy(raw, device=cuda, row_offset=row_offset)
    sc, idx = store.search(torch.from_numpy(q).to(cuda), k)
    sc, idx = sc.cpu().numpy(), idx.cpu().numpy()
    emb = so.normalise_store(raw.astype(np.float32))
    ref_i, ref_s = so.batched_topk(emb, q.astype(np.float32), k)
    assert sc.shape == ref_s.shape and idx.shape == ref_i.shape
    assert np.all(np.diff(sc, axis=1) <= 0), "scores not descending"
    for j in range(q.shape[0]):
        assert so.topk_equal_up_to_ties(idx[j] - row_offset, sc[j], ref_i[j], ref_s[j], EPS), f"query {j}"
        assert len(set(idx[j].tolist())) == idx.shape[1], "duplicate row in a result list"
    return sc, idx


@pytest.mark.parametrize("name", STORES)
@pytest.mark.parametrize("k", [1, 10, 30, 100, 1000])
def test_golden_reference_stores(pkg, cuda, name, k):
    z = np.load(GOLDEN)
    raw, q = z[f"{name}/raw"], z[f"{name}/queries"]
    store = pkg.search.FrameStore.from_numpy(raw, device=cuda)
    sc, idx = store.search(torch.from_numpy(q).to(cuda), k)
    sc, idx = sc.cpu().numpy(), idx.cpu().numpy()
    gi, gs = z[f"{name}/top{k}/idx"], z[f"{name}/top{k}/score"]
    assert idx.shape == gi.shape
    for j in range(q.shape[0]):
        assert so.topk_equal_up_to_ties(idx[j], sc[j], gi[j], gs[j], EPS if raw.dtype == np.float32 else 2 * EPS)
    # row i <-> metadata item i: the names a drop-in returns are the reference's names
    frames = z[f"{name}/frames"]
    assert all(frames[i] in frames for i in idx[0])


@pytest.mark.parametrize("N,Q,k", [(1, 1, 1), (127, 2, 5), (128, 1, 128), (129, 3, 10), (1000, 3, 100),
                                   (4096, 8, 10), (50_000, 130, 100), (20_000, 5, 300), (33_333, 257, 7),
                                   (300_000, 1, 100), (5_000, 2, 512), (70_001, 129, 1)])
def test_synthetic_vs_oracle(pkg, cuda, N, Q, k):
    rng = np.random.default_rng(N * 7 + Q)
    raw = rng.standard_normal((N, 512)).astype(np.float32)
    q = _unit(rng.standard_normal((Q, 512)).astype(np.float32))
    _check(pkg, cuda, raw, q, k)


@pytest.mark.parametrize("D", [64, 128, 256, 448, 640, 768])
def test_other_embedding_widths(pkg, cuda, D):
    rng = np.random.default_rng(D)
    raw = rng.standard_normal((3000, D)).astype(np.float32)
    q = _unit(rng.standard_normal((4, D)).astype(np.float32))
    _check(pkg, cuda, raw, q, 20)


def test_vit_l14_width_large_store_and_query_batch(pkg, cuda):
    """D = 768 (ViT-L/14 embeddings, BASELINE config 5): the shallower 2-stage row ring must still be exact, on the
    one-CTA path (Q <= 128), on the CTA-pair path (Q > 128) and with the sample pre-pass (N >= 32768)."""
    rng = np.random.default_rng(768)
    raw = rng.standard_normal((40_000, 768)).astype(np.float32)
    _check(pkg, cuda, raw, _unit(rng.standard_normal((3, 768)).astype(np.float32)), 100)
    _check(pkg, cuda, raw[:9_000], _unit(rng.standard_normal((200, 768)).astype(np.float32)), 10)


def test_fp16_unnormalised_store(pkg, cuda):
    # image_embeddings.npy of the reference is fp16 with norms 8-11 (legacy ingest, embedding.py:55)
    rng = np.random.default_rng(5)
    raw = (rng.standard_normal((2000, 512)) * 0.45).astype(np.float16)
    q = _unit(rng.standard_normal((6, 512)).astype(np.float32))
    store = pkg.search.FrameStore.from_numpy(raw, device=cuda)
    emb16 = so.normalise_store(raw)  # numpy keeps fp16
    assert np.abs(store.rows.float().cpu().numpy() - emb16.astype(np.float32)).max() <= 2.0 ** -11
    sc, idx = store.search(torch.from_numpy(q).to(cuda), 50)
    ref_i, ref_s = so.batched_topk(emb16, q, 50)
    for j in range(6):
        assert so.topk_equal_up_to_ties(idx[j].cpu().numpy(), sc[j].cpu().numpy(), ref_i[j], ref_s[j], 2 * EPS)


def test_k_larger_than_store_returns_all_sorted(pkg, cuda):
    rng = np.random.default_rng(11)
    raw = rng.standard_normal((37, 512)).astype(np.float32)
    q = _unit(rng.standard_normal((2, 512)).astype(np.float32))
    sc, idx = _check(pkg, cuda, raw, q, 100)  # embedding_service.py:317-318: len(sim) <= top_k -> all rows
    assert sc.shape == (2, 37)
    assert sorted(idx[0].tolist()) == list(range(37))


def test_planted_duplicates_and_exact_ties(pkg, cuda):
    rng = np.random.default_rng(13)
    raw = rng.standard_normal((10_000, 512)).astype(np.float32)
    q = _unit(rng.standard_normal((4, 512)).astype(np.float32))
    for j in range(4):  # plant each query three times: exact ties at the top
        for r in (100 + j, 5000 + j, 9000 + j):
            raw[r] = q[j] * 3.0
    sc, idx = _check(pkg, cuda, raw, q, 10)
    for j in range(4):
        assert sorted(idx[j, :3].tolist()) == [100 + j, 5000 + j, 9000 + j]
        assert idx[j, :3].tolist() == sorted(idx[j, :3].tolist())  # ties broken by ascending row id
        assert np.all(np.abs(sc[j, :3] - 1.0) < EPS)


def test_zero_rows_stay_zero(pkg, cuda):
    rng = np.random.default_rng(17)
    raw = rng.standard_normal((500, 512)).astype(np.float32)
    raw[7] = 0.0  # undecodable frame (embedding_service.py:477-480); the reference would emit NaN
    store = pkg.search.FrameStore.from_numpy(raw, device=cuda)
    rows = store.rows.float().cpu().numpy()
    assert np.all(rows[7] == 0.0) and np.isfinite(rows).all()
    q = _unit(rng.standard_normal((1, 512)).astype(np.float32))
    sc, idx = store.search(torch.from_numpy(q).to(cuda), 500)
    assert np.isfinite(sc.cpu().numpy()).all()


def test_row_offset_gives_global_ids(pkg, cuda):
    rng = np.random.default_rng(19)
    raw = rng.standard_normal((3000, 512)).astype(np.float32)
    q = _unit(rng.standard_normal((3, 512)).astype(np.float32))
    _, idx = _check(pkg, cuda, raw, q, 10, row_offset=5_000_000_000)
    assert idx.min() >= 5_000_000_000


def test_empty_inputs(pkg, cuda):
    store = pkg.search.FrameStore.from_numpy(np.zeros((0, 512), np.float32), device=cuda)
    sc, idx = store.search(torch.zeros((2, 512), device=cuda), 10)
    assert sc.shape == (2, 0) and idx.shape == (2, 0)
    store = pkg.search.FrameStore.from_numpy(np.ones((5, 512), np.float32), device=cuda)
    sc, idx = store.search(torch.zeros((0, 512), device=cuda), 3)
    assert sc.shape == (0, 3)


def test_rescore_matches_query_confidence(pkg, cuda):
    rng = np.random.default_rng(23)
    raw = rng.standard_normal((4000, 512)).astype(np.float32)
    q = _unit(rng.standard_normal((1, 512)).astype(np.float32))
    store = pkg.search.FrameStore.from_numpy(raw, device=cuda)
    rows = [0, 17, 3999, 128, 2048]
    out = store.rescore(q[0], rows).cpu().numpy()
    emb = so.normalise_store(raw)
    for r, o in zip(rows, out):
        assert abs(so.query_confidence(emb, r, q) - o) < EPS
    # the re-score of a search hit equals the score the scan returned (same fp16 operands)
    sc, idx = store.search(q, 5)
    again = store.rescore(q[0], idx[0]).cpu().numpy()
    assert np.allclose(again, sc[0].cpu().numpy(), atol=2e-6)


def test_merge_matches_numpy(pkg, cuda):
    rng = np.random.default_rng(29)
    P, Q, k = 8, 33, 100
    sc = rng.standard_normal((P, Q, k)).astype(np.float32)
    sc[3, :, 90:] = -np.inf  # a short shard pads with -inf / -1
    idx = rng.permutation(P * Q * k).reshape(P, Q, k).astype(np.int64)
    idx[3, :, 90:] = -1
    ms, mi = pkg.search.merge_topk(torch.from_numpy(sc).to(cuda), torch.from_numpy(idx).to(cuda))
    ms, mi = ms.cpu().numpy(), mi.cpu().numpy()
    for j in range(Q):
        flat_s, flat_i = sc[:, j].ravel(), idx[:, j].ravel()
        order = np.lexsort((flat_i, -flat_s))[:k]
        assert np.array_equal(ms[j], flat_s[order])
        assert np.array_equal(mi[j], flat_i[order])


def test_sharded_equals_unsharded_on_one_gpu(pkg, cuda):
    """Row partition + per-shard top-k + merge == one scan over the whole store, emulated on one GPU."""
    rng = np.random.default_rng(31)
    N, Q, k, P = 40_000, 9, 100, 4
    raw = rng.standard_normal((N, 512)).astype(np.float32)
    q = torch.from_numpy(_unit(rng.standard_normal((Q, 512)).astype(np.float32))).to(cuda)
    whole = pkg.search.FrameStore.from_numpy(raw, device=cuda)
    ws, wi = whole.search(q, k)
    parts_s, parts_i = [], []
    for r in range(P):
        lo, hi = pkg.search.shard_bounds(N, P, r)
        s, i = pkg.search.FrameStore.from_numpy(raw[lo:hi], device=cuda, row_offset=lo).search(q, k)
        parts_s.append(s)
        parts_i.append(i)
    ms, mi = pkg.search.merge_topk(torch.stack(parts_s), torch.stack(parts_i))
    assert torch.equal(ms, ws) and torch.equal(mi, wi)  # bit-exact: same operands, same fp32 accumulation order


@pytest.mark.parametrize("N", [10_000_000])
def test_full_size_properties(pkg, cuda, N):
    """BASELINE config 3 store size (10 M x 512 fp16 = 10.24 GB): planted rows are found, lists are sorted and
    duplicate-free, and a chunked torch fp32 scan of the same fp16 store agrees on the top-100."""
    g = torch.Generator(device=cuda).manual_seed(2)
    rows = torch.empty((N, 512), dtype=torch.float16, device=cuda)
    for r0 in range(0, N, 1_000_000):
        x = torch.randn((1_000_000, 512), generator=g, device=cuda)
        rows[r0:r0 + 1_000_000] = (x / x.norm(dim=-1, keepdim=True)).half()
    Q, k = 16, 100
    q = torch.randn((Q, 512), generator=g, device=cuda)
    q = (q / q.norm(dim=-1, keepdim=True)).half()
    planted = torch.randint(0, N, (Q,), generator=g, device=cuda)
    rows[planted] = q
    store = pkg.search.FrameStore(rows)
    sc, idx = store.search(q, k)
    assert torch.equal(idx[:, 0], planted)
    assert torch.all((sc[:, 0] - 1.0).abs() < 2e-3)
    assert torch.all(sc[:, 1:] <= sc[:, :-1])
    assert all(len(set(r.tolist())) == k for r in idx.cpu())
    best_s = torch.full((Q, k), -float("inf"), device=cuda)
    best_i = torch.full((Q, k), -1, dtype=torch.int64, device=cuda)
    for r0 in range(0, N, 1_000_000):
        s = q.float() @ rows[r0:r0 + 1_000_000].float().t()
        cs, ci = torch.cat([best_s, s], 1).topk(k, dim=1)
        all_i = torch.cat([best_i, torch.arange(r0, r0 + s.shape[1], device=cuda).expand(Q, -1)], 1)
        best_s, best_i = cs, all_i.gather(1, ci)
    assert torch.allclose(sc, best_s, atol=1e-5)
    for j in range(Q):
        assert so.topk_equal_up_to_ties(idx[j].cpu().numpy(), sc[j].cpu().numpy(), best_i[j].cpu().numpy(),
                                        best_s[j].cpu().numpy(), 1e-5)


# ---- any top_k (the reference sorts all N scores: embedding_service.py:317-320) ------------------------------------
@pytest.mark.parametrize("N,Q,k", [(5_000, 3, 2_000), (5_000, 11, 1_025), (1_500, 2, 1_500), (40_000, 9, 4_097)])
def test_large_k_matches_oracle(pkg, cuda, N, Q, k):
    """k > 1024 leaves the fused kernel (score pass + radix select over global memory); same contract."""
    rng = np.random.default_rng(N + k)
    raw = rng.standard_normal((N, 512)).astype(np.float32)
    raw[N // 2] = raw[N // 3]  # an exact tie somewhere in the middle: broken by ascending row
    q = _unit(rng.standard_normal((Q, 512)).astype(np.float32))
    sc, idx = _check(pkg, cuda, raw, q, k, row_offset=7_000)
    assert sc.shape == (Q, k)


def test_large_k_through_the_service_contract(pkg, cuda):
    """top_k * 3 > 1024 used to raise inside FrameStore.search and come back as [] from the service."""
    rng = np.random.default_rng(41)
    raw = rng.standard_normal((3_000, 512)).astype(np.float32)
    q = _unit(rng.standard_normal((1, 512)).astype(np.float32))
    store = pkg.search.FrameStore.from_numpy(raw, device=cuda)
    sc, idx = store.search(torch.from_numpy(q).to(cuda), 400 * 3)
    ref_i, ref_s = so.batched_topk(so.normalise_store(raw), q, 1200)
    assert so.topk_equal_up_to_ties(idx[0].cpu().numpy(), sc[0].cpu().numpy(), ref_i[0], ref_s[0], EPS)


def test_large_k_merge_matches_numpy(pkg, cuda):
    rng = np.random.default_rng(43)
    P, Q, k = 3, 5, 1_500
    sc = rng.standard_normal((P, Q, k)).astype(np.float32)
    sc[1, :, 1_400:] = -np.inf
    idx = rng.permutation(P * Q * k).reshape(P, Q, k).astype(np.int64)
    idx[1, :, 1_400:] = -1
    ms, mi = pkg.search.merge_topk(torch.from_numpy(sc).to(cuda), torch.from_numpy(idx).to(cuda))
    ms, mi = ms.cpu().numpy(), mi.cpu().numpy()
    for j in range(Q):
        flat_s, flat_i = sc[:, j].ravel(), idx[:, j].ravel()
        order = np.lexsort((flat_i, -flat_s))[:k]
        assert np.array_equal(ms[j], flat_s[order]) and np.array_equal(mi[j], flat_i[order])


# ---- stores that are not iid: the sample pass must stay a valid (and useful) lower bound --------------------------
def _torch_reference_topk(rows_fp16, q_fp16, k, chunk=500_000):
    """Chunked fp32 scan of the same fp16 operands (checker for stores too large for the numpy oracle)."""
    Q, N = q_fp16.shape[0], rows_fp16.shape[0]
    dev = rows_fp16.device
    best_s = torch.full((Q, k), -float("inf"), device=dev)
    best_i = torch.full((Q, k), -1, dtype=torch.int64, device=dev)
    for r0 in range(0, N, chunk):
        s = q_fp16.float() @ rows_fp16[r0:r0 + chunk].float().t()
        cs, ci = torch.cat([best_s, s], 1).topk(k, dim=1)
        all_i = torch.cat([best_i, torch.arange(r0, r0 + s.shape[1], device=dev).expand(Q, -1)], 1)
        best_s, best_i = cs, all_i.gather(1, ci)
    return best_s, best_i


@pytest.mark.parametrize("layout", ["clustered", "ascending", "descending", "mass_ties"])
@pytest.mark.parametrize("Q", [4, 300])
def test_structured_stores(pkg, cuda, layout, Q):
    """Real stores are laid out video by video in temporal order (near-duplicate neighbours, drifting content), not
    iid: clustered rows, rows sorted by their score against the query direction (both ways: the threshold either
    never settles or the best rows come last), and mass exact ties (which overflow the select kernel's shared-memory
    staging and take its global-memory passes).  Checked against a chunked fp32 scan of the same fp16 operands."""
    g = torch.Generator(device=cuda).manual_seed({"clustered": 1, "ascending": 2, "descending": 3, "mass_ties": 4}[layout] * 1000 + Q)
    N, k = 200_000, 100
    base = torch.randn((Q, 512), generator=g, device=cuda)
    base = base / base.norm(dim=-1, keepdim=True)
    if layout == "clustered":       # 200 "videos" of 1000 near-duplicate frames each
        centres = torch.randn((200, 512), generator=g, device=cuda)
        rows = centres.repeat_interleave(1000, 0) + 0.05 * torch.randn((N, 512), generator=g, device=cuda)
    elif layout in ("ascending", "descending"):
        rows = torch.randn((N, 512), generator=g, device=cuda)
        score = (rows / rows.norm(dim=-1, keepdim=True)) @ base[0]
        rows = rows[torch.argsort(score, descending=(layout == "descending"))]
    else:                           # 60 % of the store is one of 3 vectors: thousands of exact ties per query
        rows = torch.randn((N, 512), generator=g, device=cuda)
        pick = torch.randint(0, 5, (N,), generator=g, device=cuda)
        for j in range(3):
            rows[pick == j] = base[j % Q] * 2.0
    store = pkg.search.FrameStore.from_tensor(rows, device=cuda)
    q = base.half()
    sc, idx = store.search(q, k)
    ref_s, ref_i = _torch_reference_topk(store.rows, q, k)
    assert torch.allclose(sc, ref_s, atol=1e-5)
    for j in range(Q):
        assert so.topk_equal_up_to_ties(idx[j].cpu().numpy(), sc[j].cpu().numpy(), ref_i[j].cpu().numpy(),
                                        ref_s[j].cpu().numpy(), 1e-5), (layout, j)
        assert len(set(idx[j].tolist())) == k
    if layout == "mass_ties":       # ties broken by ascending row id, exactly
        s0 = store.rows.float() @ q[0].float()
        tied = (s0 == s0.max()).nonzero().flatten()
        assert tied.numel() > 10 * k and torch.equal(idx[0], tied[:k])


def test_full_size_cta_pair_path(pkg, cuda):
    """The configuration bench.py times (BASELINE config 3): 10 M x 512 fp16 rows, a query batch large enough for the
    CTA-pair kernel (topk_scan_kernel<16, 2>), the 307 200-row strided sample pass and the multi-slice plan.  64 random
    queries of the batch are compared with a chunked fp32 scan of the same operands; every list is checked for order,
    range and duplicates; planted rows must come first."""
    N, Q, k = 10_000_000, 1_024, 100
    g = torch.Generator(device=cuda).manual_seed(2)
    rows = torch.empty((N, 512), dtype=torch.float16, device=cuda)
    for r0 in range(0, N, 1_000_000):
        x = torch.randn((1_000_000, 512), generator=g, device=cuda)
        rows[r0:r0 + 1_000_000] = (x / x.norm(dim=-1, keepdim=True)).half()
        del x
    q = torch.randn((Q, 512), generator=g, device=cuda)
    q = (q / q.norm(dim=-1, keepdim=True)).half()
    planted = torch.randperm(N, generator=g, device=cuda)[:Q]
    rows[planted] = q
    store = pkg.search.FrameStore(rows)
    sc, idx = store.search(q, k)
    assert torch.equal(idx[:, 0], planted)
    assert torch.all(sc[:, 1:] <= sc[:, :-1]) and int(idx.min()) >= 0 and int(idx.max()) < N
    srt = idx.sort(dim=1).values
    assert bool((srt[:, 1:] != srt[:, :-1]).all()), "duplicate row in a result list"
    pick = torch.randperm(Q, generator=g, device=cuda)[:64]
    ref_s, ref_i = _torch_reference_topk(rows, q[pick], k, chunk=1_000_000)
    assert torch.allclose(sc[pick], ref_s, atol=1e-5)
    for j, p in enumerate(pick.tolist()):
        assert so.topk_equal_up_to_ties(idx[p].cpu().numpy(), sc[p].cpu().numpy(), ref_i[j].cpu().numpy(),
                                        ref_s[j].cpu().numpy(), 1e-5), p


@pytest.mark.parametrize("N,Q,k", [(400_000, 19_201, 10),    # 76 query tiles of 256: more tiles than CTA pairs
                                   (1_200_000, 3_329, 100),  # 14 tiles (13 full + 1 of one query), many slices, ring wraps
                                   (90_000, 5_000, 37)])      # 20 tiles over a handful of slices: more units than items at the end
def test_many_query_tiles_dynamic_items(pkg, cuda, N, Q, k):
    """The multi-tile plans hand their (query tile, slice) items out dynamically (one atomic counter per call, tickets
    published by the leader CTA's producer thread).  Whatever order the units take them in, the lists must equal a
    chunked fp32 scan of the same operands (up to ties inside 1e-5) for sampled queries of every query tile, planted
    rows must come first, and two calls on the same workspace must agree bit for bit."""
    g = torch.Generator(device=cuda).manual_seed(N + Q)
    rows = torch.randn((N, 512), generator=g, device=cuda)
    rows = (rows / rows.norm(dim=-1, keepdim=True)).half()
    q = torch.randn((Q, 512), generator=g, device=cuda)
    q = (q / q.norm(dim=-1, keepdim=True)).half()
    planted = torch.randperm(N, generator=g, device=cuda)[:Q]
    rows[planted] = q
    store = pkg.search.FrameStore(rows)
    sc, idx = store.search(q, k)
    sc2, idx2 = store.search(q, k)
    assert torch.equal(idx, idx2) and torch.equal(sc, sc2)
    assert torch.equal(idx[:, 0], planted)
    assert torch.all(sc[:, 1:] <= sc[:, :-1]) and int(idx.min()) >= 0 and int(idx.max()) < N
    pick = torch.arange(0, Q, max(1, Q // 96), device=cuda)      # a few queries of every 256-query tile
    ref_s, ref_i = _torch_reference_topk(rows, q[pick], k, chunk=200_000)
    assert torch.allclose(sc[pick], ref_s, atol=1e-5)
    for j, p in enumerate(pick.tolist()):
        assert so.topk_equal_up_to_ties(idx[p].cpu().numpy(), sc[p].cpu().numpy(), ref_i[j].cpu().numpy(),
                                        ref_s[j].cpu().numpy(), 1e-5), p
