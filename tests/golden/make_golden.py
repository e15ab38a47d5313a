"""Generates tests/golden/search_golden.npz from the reference's own stored fixtures.

Run in the authoring container only (reads /root/reference, which does not exist on the GPU box):
    python tests/golden/make_golden.py
For every store under Backend/embedding it evaluates the LITERAL numpy expressions of
Backend/services/embedding_service.py:210,314,317-320 (not the oracle functions) and
stores inputs + outputs, so tests can pin oracle/search_oracle.py and the CUDA path.
"""
import json
import os
import numpy as np

REF = "/root/reference/Backend"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "search_golden.npz")
STORES = ["video_test_3", "video_test_4", "Violence", "Violence_2"]
TOPKS = [1, 10, 30, 100, 1000]


def main():
    rng = np.random.default_rng(20261104)
    blob = {}
    mapping = json.load(open(f"{REF}/metadata/video_mapping.json"))
    blob["mapping_keys"] = np.array(list(mapping.keys()))
    for name in STORES:
        raw = np.load(f"{REF}/embedding/{name}_embeddings.npy")
        meta = json.load(open(f"{REF}/metadata/{name}_metadata.json"))
        frames = [os.path.basename(s["filepath"]) for s in meta]      # data_service.py:48
        assert len(frames) == raw.shape[0]
        emb = raw / np.linalg.norm(raw, axis=-1, keepdims=True)       # :210
        n, d = emb.shape
        # queries: 4 self rows (fp32, as get_text_features returns on CPU) + 4 random unit vectors
        self_rows = rng.choice(n, 4, replace=False)
        q_self = emb[self_rows].astype(np.float32)
        q_rand = rng.standard_normal((4, d)).astype(np.float32)
        q_rand /= np.linalg.norm(q_rand, axis=-1, keepdims=True)
        queries = np.concatenate([q_self, q_rand], 0)
        blob[f"{name}/raw"] = raw
        blob[f"{name}/frames"] = np.array(frames)
        blob[f"{name}/queries"] = queries
        blob[f"{name}/self_rows"] = self_rows
        for k in TOPKS:
            idxs, scs = [], []
            for j in range(queries.shape[0]):
                tf = queries[j:j + 1]
                similarities = np.dot(emb, tf.T).flatten()            # :314
                if len(similarities) <= k:                            # :317-320
                    top = np.argsort(similarities)[::-1]
                else:
                    top = np.argsort(similarities)[::-1][:k]
                idxs.append(top.astype(np.int64))
                scs.append(similarities[top].astype(np.float32))
            blob[f"{name}/top{k}/idx"] = np.stack(idxs)
            blob[f"{name}/top{k}/score"] = np.stack(scs)
    np.savez_compressed(OUT, **blob)
    print("wrote", OUT, os.path.getsize(OUT))


if __name__ == "__main__":
    main()
