"""Scene-cut key-frame extraction: the call surface of Backend/segment_video.py with the
per-frame arithmetic of PySceneDetect's ContentDetector on the GPU (`evr_scene_deltas`, csrc/scenecut.cu).

    extract_frames_from_video(video_path, output_dir, threshold=20.0)      # segment_video.py:6, same signature

`scenedetect` is an un-pinned dependency of the reference (absent here); its published detector is restated
(0.5.x line, the one with the `VideoManager` API the reference uses): score of frame n = mean over H, S, V of the
mean absolute difference to frame n-1 in OpenCV's 8-bit HSV; a cut at frame n needs score >= threshold and
min_scene_len (15) frames since the previous cut; scenes are the intervals between cuts; the saved key-frame is the
middle frame of each scene, written as `<frameidx>.jpg` (segment_video.py:15-24).  Video decoding stays with
cv2.VideoCapture on the host, as in the reference (no NVDEC binding in this image)."""
import os

import numpy as np
import torch

from ._lib import check, current_stream_ptr, lib, ptr

HSV_SHIFT = 12


def _hsv_tables():
    """OpenCV color_hsv: sdiv_table[i] = cvRound((255 << 12) / i), hdiv_table180[i] = cvRound((180 << 12) / (6 i))."""
    i = np.arange(1, 256, dtype=np.float64)
    sdiv = np.zeros(256, dtype=np.int32)
    hdiv = np.zeros(256, dtype=np.int32)
    sdiv[1:] = np.rint((255 << HSV_SHIFT) / (1.0 * i)).astype(np.int32)
    hdiv[1:] = np.rint((180 << HSV_SHIFT) / (6.0 * i)).astype(np.int32)
    return sdiv, hdiv


class ContentDetector:
    """Streaming detector: feed decoded frames in chunks with `process`, read `cuts`, `scene_list()`,
    `middle_frames()`.  `order`: "bgr" (cv2 frames, what the reference feeds) or "rgb"."""

    def __init__(self, threshold=27.0, min_scene_len=15, downscale=1, device="cuda", order="bgr"):
        self.threshold = float(threshold)
        self.min_scene_len = int(min_scene_len)
        self.downscale = int(downscale)
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("ContentDetector runs on a CUDA device only")
        if order not in ("bgr", "rgb"):
            raise ValueError("order must be 'bgr' or 'rgb'")
        self.bgr = 1 if order == "bgr" else 0
        sdiv, hdiv = _hsv_tables()
        self._sdiv = torch.from_numpy(sdiv).to(self.device)
        self._hdiv = torch.from_numpy(hdiv).to(self.device)
        self._last = None          # last frame of the previous chunk (device)
        self.num_frames = 0
        self.scores = []           # scores[i] belongs to frame i + 1
        self.cuts = []
        self._last_cut = None

    @torch.no_grad()
    def delta_sums(self, frames):
        """uint8 [T, H, W, 3] on the device -> int64 [T-1, 3] sums of |dH|, |dS|, |dV| (exact integers)."""
        if frames.dtype != torch.uint8 or frames.dim() != 4 or frames.shape[-1] != 3 or not frames.is_cuda:
            raise ValueError(f"frames must be a CUDA uint8 [T, H, W, 3] tensor, got {frames.dtype} {tuple(frames.shape)}")
        frames = frames.contiguous()
        T, H, W, _ = frames.shape
        sums = torch.empty((max(T - 1, 0), 3), dtype=torch.int64, device=self.device)
        done = 0
        with torch.cuda.device(self.device):
            while done < T - 1:  # grid.y limit: 65535 pairs per launch
                n = min(T - 1 - done, 65535)
                check(lib.evr_scene_deltas(ptr(frames[done:done + n + 1]), n + 1, H, W, self.bgr, self.downscale,
                                           ptr(self._sdiv), ptr(self._hdiv), ptr(sums[done:done + n]),
                                           current_stream_ptr()), "evr_scene_deltas")
                done += n
        return sums

    def process(self, frames):
        """Next chunk of consecutive frames (numpy or torch, host or device).  Returns the cuts found in it."""
        if isinstance(frames, np.ndarray):
            frames = torch.from_numpy(np.ascontiguousarray(frames))
        frames = frames.to(self.device)
        if frames.shape[0] == 0:
            return []
        chunk = frames if self._last is None else torch.cat([self._last[None], frames])
        H, W = frames.shape[1], frames.shape[2]
        f = self.downscale
        pixels = ((H + f - 1) // f) * ((W + f - 1) // f)
        sums = self.delta_sums(chunk).cpu().numpy()
        scores = (sums.astype(np.float64) / float(pixels)).sum(axis=1) / 3.0  # delta_hsv_avg
        first = self.num_frames + (0 if self._last is not None else 1)
        found = []
        for i, sc in enumerate(scores):
            n = first + i
            if sc >= self.threshold and (self._last_cut is None or n - self._last_cut >= self.min_scene_len):
                found.append(n)
                self._last_cut = n
        self.scores.extend(scores.tolist())
        self.cuts.extend(found)
        self._last = frames[-1].clone()
        self.num_frames += frames.shape[0]
        return found

    def scene_list(self):
        """[(start, end)) between cuts; no cut -> no scene, as SceneManager.get_scene_list() of that line."""
        if not self.cuts:
            return []
        edges = [0] + self.cuts + [self.num_frames]
        return [(edges[i], edges[i + 1]) for i in range(len(edges) - 1)]

    def middle_frames(self):
        return [(s + e) // 2 for s, e in self.scene_list()]  # segment_video.py:15-16


def detect_scenes(frames, threshold=27.0, min_scene_len=15, downscale=1, device="cuda", order="bgr", chunk=64):
    """All-in-one over an in-memory clip [T, H, W, 3]: -> (cuts, scene_list, middle_frames)."""
    det = ContentDetector(threshold, min_scene_len, downscale, device, order)
    for lo in range(0, frames.shape[0], chunk):
        det.process(frames[lo:lo + chunk])
    return det.cuts, det.scene_list(), det.middle_frames()


def extract_frames_from_video(video_path, output_dir, threshold=20.0, device="cuda", chunk=64, downscale=1):
    """segment_video.py:6-28 with the detector on the GPU: decode with cv2.VideoCapture (BGR, as the reference),
    score on the device chunk by chunk, then seek and save the middle frame of every scene as `<frameidx>.jpg`."""
    import cv2
    os.makedirs(output_dir, exist_ok=True)
    cap = cv2.VideoCapture(video_path)
    det = ContentDetector(threshold=threshold, device=device, order="bgr", downscale=downscale)
    buf = []
    while True:
        ret, frame = cap.read()
        if ret:
            buf.append(frame)
        if buf and (not ret or len(buf) == chunk):
            det.process(np.stack(buf))
            buf = []
        if not ret:
            break
    saved = []
    for frame_idx in det.middle_frames():
        cap.set(cv2.CAP_PROP_POS_FRAMES, frame_idx)
        ret, frame = cap.read()
        if ret:
            path = os.path.join(output_dir, f"{frame_idx}.jpg")
            cv2.imwrite(path, frame)
            print(f"Saved frame: {path}")
            saved.append(path)
    cap.release()
    print("Video extraction complete.")
    return saved
