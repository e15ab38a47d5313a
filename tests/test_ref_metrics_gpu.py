"""`metrics.evaluate_retrieval` (full rankings through the fused scan on the device) against the reference's own
`evaluate_model` outputs (tests/golden/ref_metrics_golden.json).  The store is fp16 on the device, so a rank may move
where two similarities lie within 2^-10 of each other: R@K may differ by one caption / image, MRR by 0.01, the median by
one position and the mean rank by 1 % plus two single-position moves over the set (stated tolerance); on these seeded
cases most numbers come out identical."""
import json
import os

import numpy as np
import pytest

import ref_metrics_cases as cases

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_metrics_golden.json")


def golden_cases():
    with open(GOLDEN) as f:
        return json.load(f)["cases"]


@pytest.mark.parametrize("entry", golden_cases(), ids=lambda e: e["recipe"]["name"])
def test_device_metrics_match_the_references(pkg, cuda, entry):
    case = cases.build(entry["recipe"])
    got = pkg.metrics.evaluate_retrieval(case["image_features"], case["text_features"], case["caption_image"],
                                         device=cuda)
    n = {"t2i": int((case["caption_image"] >= 0).sum()),
         "i2t": len(np.unique(case["caption_image"][case["caption_image"] >= 0]))}
    for part in ("t2i", "i2t"):
        for key in ("R@1", "R@5", "R@10"):
            assert abs(got[part][key] - entry[part][key]) <= 1.0 / n[part] + 1e-12, (part, key)
        assert abs(got[part]["MRR"] - entry[part]["MRR"]) <= 0.01
        assert abs(got[part]["Median_Rank"] - entry[part]["Median_Rank"]) <= 1.0
        assert abs(got[part]["Mean_Rank"] - entry[part]["Mean_Rank"]) <= 0.01 * entry[part]["Mean_Rank"] + 2.0 / n[part]
    assert abs(got["mean"]["rsum"] - entry["mean"]["rsum"]) <= 3.0 / n["t2i"] + 3.0 / n["i2t"] + 1e-12
    assert set(got["mean"]) == set(entry["mean"])
