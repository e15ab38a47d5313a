"""Scene-cut scoring on the GPU (`evr_scene_deltas`, scenecut.ContentDetector) against oracle/scenecut_oracle.py:
integer sums bit-exact, cuts / scene list / middle frames identical, chunked streaming == whole clip."""
import importlib

import numpy as np
import pytest
import torch

from conftest import PKG
from oracle import scenecut_oracle as so

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def scenecut(pkg):
    return importlib.import_module(PKG + ".scenecut")


@pytest.mark.parametrize("T,H,W", [(6, 72, 96), (5, 37, 41), (3, 1, 1), (2, 720, 1280), (4, 5, 3), (38, 36, 64),
                                   (17, 2, 2), (5, 720, 1280)])
@pytest.mark.parametrize("order", ["bgr", "rgb"])
@pytest.mark.parametrize("downscale", [1, 2, 3])
def test_delta_sums_bit_exact(scenecut, cuda, T, H, W, order, downscale):
    rng = np.random.default_rng(T * 1000 + H + W)
    clip = rng.integers(0, 256, (T, H, W, 3), dtype=np.uint8)
    clip[1, : H // 2] = clip[0, : H // 2]                        # identical region: zero contribution
    det = scenecut.ContentDetector(device=cuda, order=order, downscale=downscale)
    got = det.delta_sums(torch.from_numpy(clip).to(cuda)).cpu().numpy()
    ref = so.channel_delta_sums(clip, order=order, downscale=downscale)
    assert got.dtype == np.int64 and np.array_equal(got, ref)


def test_every_colour_once(scenecut, cuda):
    """All 2^24 colours against black: the sums equal the oracle's (the HSV of every colour takes part)."""
    v = np.arange(256, dtype=np.uint8)
    b, g, r = np.meshgrid(v, v, v, indexing="ij")
    img = np.stack([b, g, r], -1).reshape(4096, 4096, 3)
    clip = np.stack([np.zeros_like(img), img])
    det = scenecut.ContentDetector(device=cuda)
    got = det.delta_sums(torch.from_numpy(clip).to(cuda)).cpu().numpy()
    hsv = so.bgr_to_hsv_u8(img).astype(np.int64)
    assert np.array_equal(got[0], hsv.reshape(-1, 3).sum(0))


def test_detector_matches_oracle_and_streams(scenecut, cuda):
    clip = so.synthetic_clip(90, h=72, w=96, cuts=(17, 40, 47, 70), seed=5)   # 47 is 7 frames after 40: suppressed
    sums = so.channel_delta_sums(clip)
    scores = so.content_scores(sums, 72 * 96)
    ref_cuts = so.cuts_from_scores(scores, threshold=20.0)
    assert ref_cuts == [17, 40, 70]
    cuts, scenes, mids = scenecut.detect_scenes(clip, threshold=20.0, device=cuda, chunk=90)
    assert cuts == ref_cuts and scenes == so.scene_list(ref_cuts, 90) and mids == so.middle_frames(scenes)
    for chunk in (1, 7, 64):   # streaming in chunks carries the last frame over
        det = scenecut.ContentDetector(threshold=20.0, device=cuda)
        for lo in range(0, 90, chunk):
            det.process(clip[lo:lo + chunk])
        assert det.cuts == ref_cuts and det.num_frames == 90
        assert np.array_equal(np.asarray(det.scores), scores)
    still = np.repeat(clip[:1], 20, axis=0)
    assert scenecut.detect_scenes(still, device=cuda) == ([], [], [])


def test_argument_errors(scenecut, cuda):
    det = scenecut.ContentDetector(device=cuda)
    with pytest.raises(ValueError):
        det.delta_sums(torch.zeros((2, 4, 4, 3), dtype=torch.float32, device=cuda))
    with pytest.raises(ValueError):
        det.delta_sums(torch.zeros((2, 4, 4, 3), dtype=torch.uint8))
    with pytest.raises(RuntimeError):
        scenecut.ContentDetector(device="cpu")
    assert det.delta_sums(torch.zeros((1, 4, 4, 3), dtype=torch.uint8, device=cuda)).shape == (0, 3)


def test_video_file_round_trip(scenecut, cuda, tmp_path):
    """extract_frames_from_video (segment_video.py:6): a clip written with cv2, decoded again with cv2, scored on
    the GPU; the saved key-frames are the middle frames of the scenes the oracle finds on the decoded frames."""
    cv2 = pytest.importorskip("cv2")
    clip = so.synthetic_clip(60, h=96, w=128, cuts=(20, 41), seed=9)
    path = str(tmp_path / "clip.avi")
    wr = cv2.VideoWriter(path, cv2.VideoWriter_fourcc(*"MJPG"), 25.0, (128, 96))
    if not wr.isOpened():
        pytest.skip("no MJPG writer in this cv2 build")
    for f in clip:
        wr.write(f)
    wr.release()
    cap = cv2.VideoCapture(path)
    decoded = []
    while True:
        ret, f = cap.read()
        if not ret:
            break
        decoded.append(f)
    cap.release()
    decoded = np.stack(decoded)
    ref_cuts = so.cuts_from_scores(so.content_scores(so.channel_delta_sums(decoded), 96 * 128), threshold=20.0)
    mids = so.middle_frames(so.scene_list(ref_cuts, len(decoded)))
    saved = scenecut.extract_frames_from_video(path, str(tmp_path / "out"), threshold=20.0, device=cuda)
    assert ref_cuts == [20, 41] and [int(s.split("/")[-1][:-4]) for s in saved] == mids
