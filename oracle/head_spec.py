"""TEST INFRASTRUCTURE ONLY (oracle). Not imported by the product package.

Shape/config description shared by the oracle restatements.  Mirrors the
switch surface consumed by the reference head path:

* ``head_grid_sizes``/``class_num``/``anchor_boxes``  -> yolov3/yolov3_decoder.py:14-45
* ``iou_thresh``/``loss_weights``/``rectified_*``/``focal_*``/``is_tiou_recall``
  -> yolov3/yolov3_loss.py:16-67, configs.py:37-72
* ``eps`` -> keras.backend.epsilon(), set to 1e-8 at run.py:26

Parity status: **parity unpinned** for decode/assign/loss (the reference ships no
tests or golden vectors and TensorFlow is not installable here); post-process is
pinned against the reference's own NumPy code (see oracle/ref_postprocess.py).
"""
from dataclasses import dataclass, field
from typing import List, Sequence, Tuple

import numpy as np

# (16, 4, 1) raster-consecutive head cells folded into one /32 row:
# keras Reshape([H32, W32, -1]) at yolov3/yolov3_detector.py:80-85.
HEAD_FOLD = (16, 4, 1)

# configs.py:37-41 verbatim (normalised (w, h); 3 / 2 / 3 anchors).
SHIPPED_ANCHORS = [
    [(0.06618181818181816, 0.1025177510694752), (0.18544278606965178, 0.13160367921287464),
     (0.13, 0.32733333333333337)],
    [(0.13, 0.32733333333333337), (0.303806787732042, 0.34370030784316496)],
    [(0.303806787732042, 0.34370030784316496), (0.4667050847457627, 0.5281262429095761),
     (0.7906945888923907, 0.7888860433597275)],
]
# configs.py:52
SHIPPED_LOSS_WEIGHTS = [(5, 5, 0.05, 3, 1), (8, 8, 0.05, 2, 1), (10, 10, 0.05, 2, 1)]

# Standard COCO YOLOv3 priors in pixels at 416 (benchmark choice).
COCO_ANCHORS_416 = [
    [(10, 13), (16, 30), (33, 23)],
    [(30, 61), (62, 45), (59, 119)],
    [(116, 90), (156, 198), (373, 326)],
]


def coco_anchors():
    return [[(w / 416.0, h / 416.0) for (w, h) in head] for head in COCO_ANCHORS_416]


@dataclass
class HeadSpec:
    head_grid_sizes: Sequence[Tuple[int, int]]          # [(H8,W8),(H16,W16),(H32,W32)]
    class_num: int
    anchor_boxes: Sequence[Sequence[Tuple[float, float]]]
    iou_thresh: float = 0.8
    loss_weights: Sequence[Sequence[float]] = field(default_factory=lambda: SHIPPED_LOSS_WEIGHTS)
    rectified_loss_weight: Sequence[float] = (0.01, 0.01, 0.01)   # yolov3_loss.py:63-64
    is_focal_loss: bool = False
    focal_alpha: float = 0.25
    focal_gamma: float = 2.0
    is_tiou_recall: bool = False
    eps: float = 1e-8                                  # run.py:26

    def __post_init__(self):
        self.head_grid_sizes = [(int(h), int(w)) for (h, w) in self.head_grid_sizes]
        (h8, w8), (h16, w16), (h32, w32) = self.head_grid_sizes
        if not (h8 == 4 * h32 and w8 == 4 * w32 and h16 == 2 * h32 and w16 == 2 * w32):
            raise ValueError("head grids must be /8,/16,/32 of an input that is a multiple of 32")
        self.box_num = [len(a) for a in self.anchor_boxes]
        self.box_len = 5 + int(self.class_num)
        self.head_channels = [HEAD_FOLD[h] * self.box_num[h] * self.box_len for h in range(3)]
        self.chan_base = [0, self.head_channels[0], self.head_channels[0] + self.head_channels[1]]
        self.ctot = sum(self.head_channels)
        self.rows32 = h32 * w32
        self.anchors_per_image = sum(
            self.head_grid_sizes[h][0] * self.head_grid_sizes[h][1] * self.box_num[h] for h in range(3))

    @classmethod
    def from_input(cls, height, width, class_num, anchor_boxes, **kw):
        grids = [(height // s, width // s) for s in (8, 16, 32)]
        return cls(grids, class_num, anchor_boxes, **kw)

    def anchors_grid(self, dtype=np.float32):
        """anchor priors in grid units: fl(a) * fl([W_h, H_h])  (yolov3_decoder.py:35-40)."""
        out = []
        for h in range(3):
            H, W = self.head_grid_sizes[h]
            a = np.asarray(self.anchor_boxes[h], dtype=np.float32).astype(dtype)
            out.append(a * np.asarray([W, H], dtype=dtype))
        return out
