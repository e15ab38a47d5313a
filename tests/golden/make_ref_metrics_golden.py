"""Golden vectors for the retrieval metrics, produced by the reference ITSELF.

Runs the reference's unmodified `ModelComparison.evaluate_model`
(/root/reference/Backend/content/Test_compare_model/compare_models.py:908-1100) -- its source is cut out of the file with
`ast` and executed as it stands -- on seeded feature matrices: `self.process_image_batch` / `self.process_text_batch`
hand back rows of those matrices instead of running a model, the `DataLoader` is a plain batcher, and everything from the
similarity matrix on (np.dot, the argsort ranks of both directions, R@1/5/10, MRR, median / mean rank, the mean of both
directions and rsum) is the reference's own code.  The module itself cannot be imported here (it needs clip, open_clip,
matplotlib, torchvision at import time), which is why the method is lifted out instead.

The cases are stored as recipes (seed, sizes, noise) -- tests/ref_metrics_cases.py rebuilds the same matrices -- together
with what the reference printed back.  Authoring container only:
    python tests/golden/make_ref_metrics_golden.py
"""
import ast
import json
import os
import sys
import textwrap
import time
from collections import defaultdict

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import ref_metrics_cases as cases  # noqa: E402

REF = "/root/reference/Backend/content/Test_compare_model/compare_models.py"


def reference_evaluate_model():
    src = open(REF, encoding="utf-8").read()
    tree = ast.parse(src)
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "ModelComparison")
    fn = next(n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name == "evaluate_model")
    code = textwrap.dedent(ast.get_source_segment(src, fn))

    class DataLoader:   # the reference only iterates it
        def __init__(self, dataset, batch_size=1, shuffle=False, num_workers=0, pin_memory=False):
            assert not shuffle
            self.dataset, self.batch_size = dataset, batch_size

        def __iter__(self):
            for i in range(0, len(self.dataset), self.batch_size):
                yield self.dataset[i:i + self.batch_size]

    ns = {"np": np, "torch": torch, "time": time, "defaultdict": defaultdict, "DataLoader": DataLoader,
          "tqdm": lambda it, **kw: it, "print": lambda *a, **k: None}
    exec(compile(code, REF, "exec"), ns)
    return ns["evaluate_model"]


class FakeDataset:
    def __init__(self, image_ids, captions):
        self.image_ids = image_ids
        self.all_captions = captions

    def get_image_path(self, image_id):
        return image_id


class FakeSelf:
    def __init__(self, case):
        self.case = case
        self.dataset = FakeDataset(case["image_ids"], case["captions"])
        self.args = type("A", (), {"batch_size": 7})()
        self._row = {iid: i for i, iid in enumerate(case["image_ids"])}
        self._cap = {c["caption"]: i for i, c in enumerate(case["captions"])}

    def create_image_dataset(self, image_paths, model_name):
        return list(image_paths)

    def process_image_batch(self, batch, model_name):
        return torch.from_numpy(self.case["image_features"][[self._row[p] for p in batch]])

    def process_text_batch(self, captions, model_name):
        return self.case["text_features"][[self._cap[c] for c in captions]]

    def format_metrics_table(self, metrics, title=None):
        pass

    def format_combined_metrics_table(self, *args, **kwargs):
        pass


def main():
    evaluate_model = reference_evaluate_model()
    out = []
    for recipe in cases.RECIPES:
        case = cases.build(recipe)
        res = evaluate_model(FakeSelf(case), "original_clip")
        out.append({"recipe": recipe,
                    "t2i": {k: float(v) for k, v in res["t2i"].items()},
                    "i2t": {k: float(v) for k, v in res["i2t"].items()},
                    "mean": {k: float(v) for k, v in res["mean"].items()}})
        print(recipe["name"], {k: round(float(v), 4) for k, v in res["mean"].items()})
    with open(os.path.join(HERE, "ref_metrics_golden.json"), "w") as f:
        json.dump({"source": REF + ":908-1100 (evaluate_model, executed unmodified)", "cases": out}, f, indent=1)


if __name__ == "__main__":
    main()
