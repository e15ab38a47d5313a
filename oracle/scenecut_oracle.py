"""CPU restatement (numpy) of the scene-cut scoring behind Backend/segment_video.py:6-28.

TEST INFRASTRUCTURE (checker), never the shipped path.

The reference calls PySceneDetect: `SceneManager.add_detector(ContentDetector(threshold))`, `detect_scenes`,
`get_scene_list()` and keeps the middle frame of every scene (segment_video.py:9-16).  `scenedetect` is an un-vendored,
un-pinned dependency (absent here; the `VideoManager` API dates the code to the 0.5.x line), so its published algorithm
is restated:  per frame, cv2.cvtColor(BGR2HSV) -> per channel mean |delta| against the previous frame -> the mean of
the three channel deltas is compared with `threshold`; a cut needs `min_scene_len` (15) frames since the last one.
=> PARITY UNPINNED for the detector logic.  The arithmetic underneath it IS pinned: `bgr_to_hsv_u8` restates OpenCV's
8-bit integer RGB2HSV_b (modules/imgproc/src/color_hsv.simd.hpp: 12-bit fixed-point division tables, hue range 180)
and tests/test_oracle_scenecut.py checks it against cv2.cvtColor of this container on all 2^24 colours and against the
committed fixture tests/golden/hsv_golden.npz.
"""
import numpy as np

HSV_SHIFT = 12


def hsv_tables():
    """sdiv_table[i] = round(255 * 4096 / i), hdiv_table180[i] = round(180 * 4096 / (6 i)); entry 0 is 0.
    cvRound = round half to even (np.rint)."""
    i = np.arange(1, 256, dtype=np.float64)
    sdiv = np.zeros(256, dtype=np.int32)
    hdiv = np.zeros(256, dtype=np.int32)
    sdiv[1:] = np.rint((255 << HSV_SHIFT) / (1.0 * i)).astype(np.int32)
    hdiv[1:] = np.rint((180 << HSV_SHIFT) / (6.0 * i)).astype(np.int32)
    return sdiv, hdiv


def bgr_to_hsv_u8(img, order="bgr"):
    """uint8 [..., 3] -> uint8 [..., 3] (H in 0..179, S, V), bit for bit cv2.cvtColor(img, COLOR_BGR2HSV / RGB2HSV)."""
    sdiv, hdiv = hsv_tables()
    a = img.astype(np.int32)
    b, g, r = (a[..., 0], a[..., 1], a[..., 2]) if order == "bgr" else (a[..., 2], a[..., 1], a[..., 0])
    v = np.maximum(np.maximum(b, g), r)
    vmin = np.minimum(np.minimum(b, g), r)
    diff = v - vmin
    vr = v == r
    vg = v == g
    s = (diff * sdiv[v] + (1 << (HSV_SHIFT - 1))) >> HSV_SHIFT
    h = np.where(vr, g - b, np.where(vg, b - r + 2 * diff, r - g + 4 * diff))
    h = (h * hdiv[diff] + (1 << (HSV_SHIFT - 1))) >> HSV_SHIFT
    h = h + np.where(h < 0, 180, 0)
    return np.stack([h, s, v], axis=-1).astype(np.uint8)


def channel_delta_sums(frames, order="bgr", downscale=1):
    """frames uint8 [T, H, W, 3] -> int64 [T-1, 3]: sum over the (sub-sampled) pixels of |dH|, |dS|, |dV| between
    consecutive frames (ContentDetector: np.sum(np.abs(curr - last)) on int32 planes; no hue wrap-around)."""
    f = frames[:, ::downscale, ::downscale]
    hsv = bgr_to_hsv_u8(f, order).astype(np.int64)
    return np.abs(hsv[1:] - hsv[:-1]).sum(axis=(1, 2))


def content_scores(sums, n_pixels):
    """delta_hsv_avg of ContentDetector: mean over the three channels of (sum |d| / pixels)."""
    return (sums.astype(np.float64) / float(n_pixels)).sum(axis=1) / 3.0


def cuts_from_scores(scores, threshold=27.0, min_scene_len=15, first_frame=1):
    """scores[i] belongs to frame first_frame + i (frame 0 has no predecessor).  ContentDetector.process_frame (0.5.x):
    cut at frame n when score >= threshold and (no cut yet or n - last_cut >= min_scene_len)."""
    cuts, last = [], None
    for i, sc in enumerate(scores):
        n = first_frame + i
        if sc >= threshold and (last is None or n - last >= min_scene_len):
            cuts.append(n)
            last = n
    return cuts


def scene_list(cuts, num_frames):
    """SceneManager.get_scene_list(): [(start, end)) pairs from the cut list; no cut -> no scene (0.5.x behaviour)."""
    if not cuts:
        return []
    edges = [0] + list(cuts) + [num_frames]
    return [(edges[i], edges[i + 1]) for i in range(len(edges) - 1)]


def middle_frames(scenes):
    """segment_video.py:15-16: middle_frame = (start + end) // 2."""
    return [(s + e) // 2 for s, e in scenes]


def synthetic_clip(n_frames=60, h=72, w=96, cuts=(17, 40), seed=0):
    """A clip of smooth random 'shots' with small per-frame noise and hard cuts at the given frame numbers (BGR)."""
    rng = np.random.default_rng(seed)
    frames = np.empty((n_frames, h, w, 3), dtype=np.uint8)
    edges = [0] + list(cuts) + [n_frames]
    for a, b in zip(edges[:-1], edges[1:]):
        low = rng.integers(0, 256, (h // 8, w // 8, 3)).astype(np.float64)
        shot = np.kron(low, np.ones((8, 8, 1)))
        for t in range(a, b):
            frames[t] = np.clip(shot + rng.normal(0, 2.0, shot.shape), 0, 255).astype(np.uint8)
    return frames
