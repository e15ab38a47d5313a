"""Legacy ingest entry point: `extract_and_save_embeddings_from_folder(folder_path, model_name, video_name=None)`.

Mirror of Backend/embedding.py:9-59: walk the folder (`os.walk` order, .jpg/.jpeg/.png),
preprocess each image, `encode_image`, and `np.save` the UN-normalised `[N, D]` rows in the model dtype (fp16 on
CUDA) to `<embeddings_dir>/<video>_embeddings.npy`; returns that path.  The reference loads the model on every
call (:22) and runs batch 1 (:46-49); here the model is loaded once per (name, device) and frames go through the
sm_100a tower in batches, which does not change the rows (the tower is batch-invariant: tests/test_vit_gpu.py).
The reference's hard-coded `E:\\...\\Backend\\embedding` directory (:27) becomes `embeddings_dir=` /
$EVR_EMBEDDING_DIR / ./embedding.
"""
import os
import threading
from pathlib import Path

import numpy as np
import torch

from . import clip

_models = {}
_models_lock = threading.Lock()
BATCH = 256


def _model(model_name, device, load_kwargs):
    key = (model_name, str(device))
    with _models_lock:
        if key not in _models:
            _models[key] = clip.load(model_name, device=device, **(load_kwargs or {}))
        return _models[key]


def extract_and_save_embeddings_from_folder(folder_path, model_name, video_name=None, embeddings_dir=None,
                                            load_kwargs=None):
    from PIL import Image
    if not torch.cuda.is_available():
        raise RuntimeError("no CUDA device: this package has no CPU path")
    device = "cuda"
    model, preprocess = _model(model_name, device, load_kwargs)
    embeddings_dir = embeddings_dir or os.environ.get("EVR_EMBEDDING_DIR") or os.path.join(os.getcwd(), "embedding")
    os.makedirs(embeddings_dir, exist_ok=True)
    if not video_name:
        video_name = Path(folder_path).name
    output_file = os.path.join(embeddings_dir, f"{video_name}_embeddings.npy")

    all_embeddings, pending = [], []

    def flush():
        if pending:
            with torch.no_grad():
                all_embeddings.append(model.encode_image(torch.stack(pending).to(device)).cpu().numpy())
            pending.clear()

    for root, _, files in os.walk(folder_path):
        for file in files:
            if file.lower().endswith(('.jpg', '.jpeg', '.png')):
                image = Image.open(os.path.join(root, file)).convert("RGB")  # decode errors propagate, as in :45
                pending.append(preprocess(image))
                if len(pending) == BATCH:
                    flush()
    flush()
    out = np.concatenate(all_embeddings, 0) if all_embeddings else np.array([])  # np.array([]) as :55 on no images
    np.save(output_file, out)
    print(f"Embeddings saved to {output_file}")
    return output_file
