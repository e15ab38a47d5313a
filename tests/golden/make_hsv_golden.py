"""Generates tests/golden/hsv_golden.npz with the cv2 of this container (run here; cv2 need not exist on the GPU box):
4096 seeded colours + every grey + the 6 hue sextant boundaries, BGR -> cv2.cvtColor(COLOR_BGR2HSV)."""
import os

import cv2
import numpy as np

rng = np.random.default_rng(1234)
bgr = rng.integers(0, 256, (4096, 3), dtype=np.uint8)
grey = np.repeat(np.arange(256, dtype=np.uint8)[:, None], 3, axis=1)
edge = np.array([[0, 0, 255], [0, 255, 255], [0, 255, 0], [255, 255, 0], [255, 0, 0], [255, 0, 255],
                 [1, 0, 255], [0, 1, 255], [254, 255, 255], [255, 254, 253]], dtype=np.uint8)
bgr = np.concatenate([bgr, grey, edge])[None]
hsv = cv2.cvtColor(bgr, cv2.COLOR_BGR2HSV)
np.savez_compressed(os.path.join(os.path.dirname(__file__), "hsv_golden.npz"), bgr=bgr[0], hsv=hsv[0],
                    cv2_version=cv2.__version__)
print("wrote", bgr.shape[1], "colours, cv2", cv2.__version__)
