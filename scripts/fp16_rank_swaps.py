"""How often does an fp16 store + fp16 query (fp32 accumulate, what the scan computes) order frames differently from the
reference's fp32 numpy scan (embedding_service.py:312-333)?  Measured on the reference's own five stored .npy files.
Queries: rows of the OTHER stores (cross-video image queries, the image-search path) and unit vectors halfway between
two rows of the store (near-tie generators).  CPU only; prints one JSON object.  Run in the authoring container:
    python scripts/fp16_rank_swaps.py /root/reference/Backend/embedding
"""
import glob
import json
import os
import sys

import numpy as np


def unit(x):
    return x / np.linalg.norm(x, axis=1, keepdims=True)


def main(folder, k=100, nq=256):
    rng = np.random.default_rng(0)
    stores = {os.path.basename(p): unit(np.load(p).astype(np.float32)) for p in sorted(glob.glob(folder + "/*.npy"))}
    out = {}
    for name, s in stores.items():
        others = np.concatenate([v for n, v in stores.items() if n != name and v.shape[1] == s.shape[1]])
        q1 = others[rng.choice(len(others), min(nq, len(others)), replace=False)]
        a, b = s[rng.integers(0, len(s), nq)], s[rng.integers(0, len(s), nq)]
        q = np.concatenate([q1, unit(a + b + 1e-3 * rng.standard_normal(a.shape).astype(np.float32))])
        kk = min(k, len(s))
        exact = q @ s.T
        half = q.astype(np.float16).astype(np.float32) @ s.astype(np.float16).astype(np.float32).T
        oe = np.argsort(-exact, axis=1, kind="stable")[:, :kk]
        oh = np.argsort(-half, axis=1, kind="stable")[:, :kk]
        same_set = np.mean([len(set(x) & set(y)) / kk for x, y in zip(oe, oh)])
        pos_changed = float((oe != oh).mean())
        top1 = float((oe[:, 0] == oh[:, 0]).mean())
        n1 = len(q1)   # the first n1 queries are real rows of other videos; the rest tie two rows by construction
        real = {"top1_same": float((oe[:n1, 0] == oh[:n1, 0]).mean()),
                "positions_changed": float((oe[:n1] != oh[:n1]).mean())}
        # the score gap of every swapped adjacent pair, in fp32 score units
        gaps = []
        for i in range(len(q)):
            for j in np.nonzero(oe[i] != oh[i])[0]:
                gaps.append(abs(exact[i, oe[i, j]] - exact[i, oh[i, j]]))
        out[name] = {"rows": int(len(s)), "queries": int(len(q)), "k": int(kk), "topk_set_overlap": float(same_set),
                     "positions_changed": pos_changed, "top1_same": top1, "cross_video_queries_only": real,
                     "max_score_gap_of_a_moved_row": float(max(gaps) if gaps else 0.0),
                     "max_abs_score_error": float(np.abs(exact - half).max())}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "/root/reference/Backend/embedding")
