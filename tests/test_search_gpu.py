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
