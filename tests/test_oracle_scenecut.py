"""Pins oracle/scenecut_oracle.py: the integer BGR->HSV restatement against cv2 itself (all 2^24 colours, when cv2 is
importable) and against the committed fixture; the ContentDetector restatement on a synthetic clip with known cuts."""
import os

import numpy as np
import pytest

from oracle import scenecut_oracle as sc

GOLD = os.path.join(os.path.dirname(__file__), "golden", "hsv_golden.npz")


def test_hsv_matches_committed_cv2_fixture():
    g = np.load(GOLD)
    assert np.array_equal(sc.bgr_to_hsv_u8(g["bgr"]), g["hsv"])
    assert np.array_equal(sc.bgr_to_hsv_u8(g["bgr"][:, ::-1], order="rgb"), g["hsv"])


def test_hsv_matches_cv2_on_every_colour():
    cv2 = pytest.importorskip("cv2")
    v = np.arange(256, dtype=np.uint8)
    for b in range(0, 256, 16):  # 16 slabs of 2^20 colours
        bb, g, r = np.meshgrid(v[b:b + 16], v, v, indexing="ij")
        img = np.stack([bb, g, r], -1).reshape(16 * 256, 256, 3)
        assert np.array_equal(sc.bgr_to_hsv_u8(img), cv2.cvtColor(img, cv2.COLOR_BGR2HSV)), b


def test_tables_are_opencvs():
    sdiv, hdiv = sc.hsv_tables()
    assert sdiv[0] == 0 and hdiv[0] == 0 and sdiv[255] == 4096 and sdiv[1] == 255 * 4096
    assert hdiv[1] == 122880 and hdiv[255] == round(180 * 4096 / (6 * 255))


def test_detector_finds_the_planted_cuts():
    clip = sc.synthetic_clip(60, cuts=(17, 40), seed=3)
    sums = sc.channel_delta_sums(clip)
    scores = sc.content_scores(sums, clip.shape[1] * clip.shape[2])
    assert sc.cuts_from_scores(scores, threshold=20.0) == [17, 40]
    scenes = sc.scene_list([17, 40], 60)
    assert scenes == [(0, 17), (17, 40), (40, 60)] and sc.middle_frames(scenes) == [8, 28, 50]
    assert sc.scene_list([], 60) == []
    # min_scene_len: a second spike 5 frames after a cut is ignored, one 15 frames later is taken
    s = np.zeros(40)
    s[[9, 14, 24]] = 99.0          # frames 10, 15, 25
    assert sc.cuts_from_scores(s, 27.0, 15) == [10, 25]
    # sub-sampling changes the pixel set, not the rule
    sums2 = sc.channel_delta_sums(clip, downscale=2)
    assert sums2.shape == (59, 3) and (sums2 <= sums).all()
