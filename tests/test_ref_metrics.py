"""Retrieval metrics pinned to the reference ITSELF: tests/golden/ref_metrics_golden.json holds
what the reference's unmodified `ModelComparison.evaluate_model` (compare_models.py:908-1100) returned for the seeded
feature matrices of tests/ref_metrics_cases.py (tests/golden/make_ref_metrics_golden.py).  The oracle restatement must
reproduce every number exactly (fp64 rounding of a mean aside); the GPU path is checked in test_ref_metrics_gpu.py."""
import json
import os

import numpy as np
import pytest

import ref_metrics_cases as cases
from oracle import search_oracle as so

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_metrics_golden.json")


def golden_cases():
    with open(GOLDEN) as f:
        return json.load(f)["cases"]


@pytest.mark.parametrize("entry", golden_cases(), ids=lambda e: e["recipe"]["name"])
def test_oracle_metrics_equal_the_references(entry):
    case = cases.build(entry["recipe"])
    got = so.t2i_i2t_metrics_general(case["image_features"], case["text_features"], case["caption_image"])
    for part in ("t2i", "i2t", "mean"):
        assert set(got[part]) == set(entry[part])
        for key, want in entry[part].items():
            assert got[part][key] == pytest.approx(want, rel=0, abs=1e-12), (part, key)


def test_uniform_layout_helper_agrees_with_the_general_one():
    case = cases.build(cases.RECIPES[0])       # 5 captions per image, grouped by image
    a = so.t2i_i2t_metrics(case["image_features"], case["text_features"], 5)
    b = so.t2i_i2t_metrics_general(case["image_features"], case["text_features"], case["caption_image"])
    assert a["rsum"] == pytest.approx(b["mean"]["rsum"], abs=1e-12)
    assert a["t2i"]["MRR"] == pytest.approx(b["t2i"]["MRR"], abs=1e-12)
    assert a["i2t"]["MedianRank"] == b["i2t"]["Median_Rank"]


def test_golden_covers_the_edge_cases():
    names = {e["recipe"]["name"] for e in golden_cases()}
    assert {"uneven_and_uncaptioned", "captions_of_unknown_images", "one_image", "shuffled_captions"} <= names
    orphan = cases.build(next(r for r in cases.RECIPES if r["name"] == "captions_of_unknown_images"))
    assert (orphan["caption_image"] < 0).sum() == 9
    uncaptioned = cases.build(next(r for r in cases.RECIPES if r["name"] == "uneven_and_uncaptioned"))
    assert len(np.unique(uncaptioned["caption_image"])) < len(uncaptioned["image_ids"])
