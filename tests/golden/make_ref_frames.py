"""Copies a handful of the reference's stored key-frames next to the embedding rows the reference stored for them
(Backend/static/processed_frames/<video>/<frameidx>.jpg <-> row i of Backend/embedding/<video>_embeddings.npy, where i is
the frame's position in the sorted file list, embedding_service.py:452).  These (jpg -> row) pairs are true golden
vectors of the ORIGINAL OpenAI ViT-B/32 weights; they pin `clip.encode_image` wherever the weights file
`~/.cache/clip/ViT-B-32.pt` exists (tests/test_real_weights_gpu.py; skipped without it).  Authoring container only."""
import json
import os
import shutil

import numpy as np

REF = "/root/reference/Backend"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_frames")
PICK = {"video_test_4": ["47450.jpg", "47588.jpg", "47058.jpg"], "Violence": ["4009.jpg", "96.jpg", "4777.jpg"],
        "video_test_3": ["20582.jpg", "99.jpg"]}   # smallest files: about 0.5 MB in all


def main():
    os.makedirs(OUT, exist_ok=True)
    rows, index = [], []
    for video, files in PICK.items():
        names = sorted(os.listdir(f"{REF}/static/processed_frames/{video}"))
        raw = np.load(f"{REF}/embedding/{video}_embeddings.npy")
        meta = json.load(open(f"{REF}/metadata/{video}_metadata.json", encoding="utf-8"))
        assert len(names) == raw.shape[0] == len(meta)
        for f in files:
            i = names.index(f)
            assert meta[i]["frameid"] == f      # row i <-> metadata item i <-> sorted file i
            shutil.copyfile(f"{REF}/static/processed_frames/{video}/{f}", os.path.join(OUT, f"{video}__{f}"))
            rows.append(raw[i].astype(np.float32))
            index.append({"file": f"{video}__{f}", "video": video, "row": i, "stored_dtype": str(raw.dtype)})
    np.save(os.path.join(OUT, "rows.npy"), np.stack(rows))
    json.dump(index, open(os.path.join(OUT, "index.json"), "w"), indent=1)
    print("wrote", len(index), "pairs to", OUT)


if __name__ == "__main__":
    main()
