"""Seeded retrieval-metric cases shared by tests/golden/make_ref_metrics_golden.py (which feeds them to the reference's
own evaluate_model) and tests/test_ref_metrics*.py (which feed them to the oracle and to the GPU path)."""
import numpy as np

RECIPES = [
    {"name": "flickr_like", "seed": 1, "images": 60, "dim": 48, "caps": [5, 5], "noise": 3.0},
    {"name": "hard", "seed": 2, "images": 120, "dim": 24, "caps": [5, 5], "noise": 4.0},
    {"name": "uneven_and_uncaptioned", "seed": 3, "images": 50, "dim": 32, "caps": [0, 5], "noise": 3.0},
    {"name": "shuffled_captions", "seed": 4, "images": 40, "dim": 32, "caps": [1, 4], "noise": 2.5, "shuffle": True},
    {"name": "captions_of_unknown_images", "seed": 5, "images": 45, "dim": 40, "caps": [2, 5], "noise": 3.0,
     "shuffle": True, "orphans": 9},
    {"name": "one_image", "seed": 6, "images": 1, "dim": 16, "caps": [3, 3], "noise": 0.5},
]


def _unit(x):
    return (x / np.linalg.norm(x, axis=-1, keepdims=True)).astype(np.float32)


def build(recipe):
    """-> dict(image_ids, captions=[{'caption', 'image_id'}], image_features [I, D], text_features [T, D],
    caption_image [T] (index into image_ids, -1 for a caption whose image is not in the set))."""
    rng = np.random.default_rng(recipe["seed"])
    n, d = recipe["images"], recipe["dim"]
    image_features = _unit(rng.standard_normal((n, d)))
    image_ids = [f"img_{i:04d}.jpg" for i in range(n)]
    lo, hi = recipe["caps"]
    owners = []
    for i in range(n):
        owners += [i] * int(rng.integers(lo, hi + 1))
    owners += [-1] * recipe.get("orphans", 0)
    owners = np.array(owners, dtype=np.int64)
    if recipe.get("shuffle"):
        owners = owners[rng.permutation(len(owners))]
    base = np.where(owners[:, None] >= 0, image_features[np.maximum(owners, 0)],
                    _unit(rng.standard_normal((len(owners), d))))
    text_features = _unit(base + recipe["noise"] * rng.standard_normal((len(owners), d)) / np.sqrt(d))
    captions = [{"caption": f"caption {t}", "image_id": image_ids[o] if o >= 0 else f"missing_{t}.jpg"}
                for t, o in enumerate(owners)]
    return {"image_ids": image_ids, "captions": captions, "image_features": image_features,
            "text_features": text_features, "caption_image": owners}
