"""Retrieval metrics of the reference's evaluation harness, computed through the fused scan.

Mirror of `ModelComparison.evaluate_model` in Backend/content/Test_compare_model/compare_models.py:
    similarity_matrix = np.dot(image_features, text_features.T)                     :999
    t2i: per caption, rank of its image in argsort(-similarities[:, i])              :1004-1016
    i2t: per image, best rank among its captions in argsort(-similarities[j, :])     :1046-1062
    R@1 / R@5 / R@10 / MRR / Median_Rank / Mean_Rank                                  :1019-1027, :1065-1072
    mean of both directions + rsum                                                    :1075-1084
The similarity matrix and its per-column argsort are replaced by `FrameStore.search` with k = all rows (the full
ranking of every query, on the device; stores above 1024 rows take the any-k path), so a rank is the position of the
ground-truth row in the returned list.  Ties: the scan orders equal scores by ascending row, numpy's argsort leaves
them unspecified.
"""
import numpy as np
import torch

from .search import FrameStore

RECALL_AT = (1, 5, 10)


def metrics_from_ranks(ranks):
    """compare_models.py:1019-1027 -- ranks are 1-based."""
    ranks = np.asarray(ranks)
    out = {f"R@{k}": float((ranks <= k).mean()) for k in RECALL_AT}
    out.update({"MRR": float((1.0 / ranks).mean()), "Median_Rank": float(np.median(ranks)),
                "Mean_Rank": float(np.mean(ranks))})
    return out


def full_ranking(store_rows, queries, device="cuda", query_chunk=256):
    """[Q, N] int64: row ids of `store_rows` in descending cosine for every query (both are L2-normalised here)."""
    rows, q = torch.as_tensor(store_rows), torch.as_tensor(queries)
    pad = (-rows.shape[1]) % 64        # the scan takes widths that are multiples of 64: zero columns change no dot product
    if pad:
        rows = torch.nn.functional.pad(rows, (0, pad))
        q = torch.nn.functional.pad(q, (0, pad))
    store = FrameStore.from_tensor(rows, device=device)
    out = []
    for i in range(0, q.shape[0], query_chunk):
        out.append(store.search(q[i:i + query_chunk], store.N)[1])
    return torch.cat(out) if out else torch.empty((0, store.N), dtype=torch.int64, device=store.device)


def positions_of(ranking, targets):
    """1-based position of targets[q] in ranking[q, :]."""
    hit = ranking == torch.as_tensor(targets, device=ranking.device).view(-1, 1)
    if not bool(hit.any(dim=1).all()):
        raise ValueError("a ground-truth row is missing from its ranking")
    return (hit.int().argmax(dim=1) + 1).cpu().numpy()


def evaluate_retrieval(image_features, text_features, caption_image, device="cuda"):
    """image_features [I, D], text_features [T, D] (any float dtype, normalised or not), caption_image [T]: index of the
    image each caption describes (-1: its image is not in the set) -> {"t2i": {...}, "i2t": {...},
    "mean": {..., "rsum"}, "t2i_ranks", "i2t_ranks"}."""
    caption_image = np.asarray(caption_image, dtype=np.int64)
    known = np.nonzero(caption_image >= 0)[0]     # a caption of an image outside the set is skipped (:1007) ...
    t2i_ranks = positions_of(full_ranking(image_features, torch.as_tensor(text_features)[known], device),
                             caption_image[known])
    images_with_captions = np.unique(caption_image[known])     # ... but still competes in i2t; :1052-1054 skips images without a caption
    order = full_ranking(text_features, torch.as_tensor(image_features)[images_with_captions], device)
    pos = torch.empty_like(order)
    pos.scatter_(1, order, torch.arange(order.shape[1], device=order.device).expand_as(order))   # caption -> position
    pos = pos.cpu().numpy() + 1
    i2t_ranks = np.array([pos[j, caption_image == img].min() for j, img in enumerate(images_with_captions)])
    t2i, i2t = metrics_from_ranks(t2i_ranks), metrics_from_ranks(i2t_ranks)
    mean = {k: (t2i[k] + i2t[k]) / 2 for k in t2i}
    mean["rsum"] = sum(t2i[f"R@{k}"] + i2t[f"R@{k}"] for k in RECALL_AT)
    return {"t2i": t2i, "i2t": i2t, "mean": mean, "t2i_ranks": t2i_ranks, "i2t_ranks": i2t_ranks}
