"""Synthetic head tensors and labels of the shapes BASELINE.json names.

predicts: fp32, packed (N, H/32, W/32, Ctot) layout (reference yolov3/yolov3_detector.py:80-85),
iid N(0,1) clipped to [-8, 8] so fp32 sigmoid stays clear of the saturation described in SURVEY F7.
targets:  (N, M*5) rows [x, y, w, h, cls]*M normalised to the input, -1 padded like
dataset/file_util.py:97.
"""
import numpy as np

COCO_ANCHORS_416 = [
    [(10, 13), (16, 30), (33, 23)],
    [(30, 61), (62, 45), (59, 119)],
    [(116, 90), (156, 198), (373, 326)],
]
# configs.py:37-41
SHIPPED_ANCHORS = [
    [(0.06618181818181816, 0.1025177510694752), (0.18544278606965178, 0.13160367921287464),
     (0.13, 0.32733333333333337)],
    [(0.13, 0.32733333333333337), (0.303806787732042, 0.34370030784316496)],
    [(0.303806787732042, 0.34370030784316496), (0.4667050847457627, 0.5281262429095761),
     (0.7906945888923907, 0.7888860433597275)],
]


def coco_anchors():
    return [[(w / 416.0, h / 416.0) for (w, h) in head] for head in COCO_ANCHORS_416]


def three_anchor_shipped():
    """config 5: shipped priors with the /16 head extended to three."""
    a = [list(h) for h in SHIPPED_ANCHORS]
    a[1] = a[1] + [SHIPPED_ANCHORS[2][1]]
    return a


def head_grid_sizes(height, width):
    if height % 32 or width % 32:
        raise ValueError("input size must be a multiple of 32")
    return [(height // s, width // s) for s in (8, 16, 32)]


def packed_channels(box_num, class_num):
    L = 5 + class_num
    return (16 * box_num[0] + 4 * box_num[1] + box_num[2]) * L


def make_predicts(n, height, width, box_num, class_num, seed=1234, clip=8.0):
    rng = np.random.default_rng(seed)
    shape = (n, height // 32, width // 32, packed_channels(box_num, class_num))
    p = rng.standard_normal(shape).astype(np.float32)
    return np.clip(p, -clip, clip).astype(np.float32)


def make_targets(n, class_num, m=16, seed=4321, min_valid=1, max_valid=None):
    rng = np.random.default_rng(seed)
    max_valid = m if max_valid is None else max_valid
    t = np.full((n, m, 5), -1.0, dtype=np.float32)
    for i in range(n):
        v = int(rng.integers(min_valid, max_valid + 1)) if max_valid >= min_valid else 0
        if v == 0:
            continue
        xy = rng.uniform(0.02, 0.98, size=(v, 2))
        wh = np.exp(rng.uniform(np.log(0.02), np.log(0.6), size=(v, 2)))
        # keep the box inside the image
        wh = np.minimum(wh, 2.0 * np.minimum(xy, 1.0 - xy))
        cls = rng.integers(0, max(class_num, 1), size=(v, 1))
        t[i, :v] = np.concatenate([xy, wh, cls], axis=1).astype(np.float32)
    return t.reshape(n, m * 5)


# configs.py:52 (xy, wh, noobj, obj, cls per head)
SHIPPED_LOSS_WEIGHTS = [(5, 5, 0.05, 3, 1), (8, 8, 0.05, 2, 1), (10, 10, 0.05, 2, 1)]


class ProductSpec(object):
    """The switches and derived sizes the product classes and bench.py need for one workload."""

    def __init__(self, height, width, class_num, anchor_boxes, iou_thresh, loss_weights=None, rectified_loss_weight=(1.0, 1.0, 1.0),
                 is_focal_loss=False, focal_alpha=1.0, focal_gamma=2.0, is_tiou_recall=False, eps=1e-8):
        self.head_grid_sizes = head_grid_sizes(height, width)
        self.class_num = int(class_num)
        self.anchor_boxes = anchor_boxes
        self.box_num = [len(a) for a in anchor_boxes]
        self.box_len = 5 + self.class_num
        fold = (16, 4, 1)
        self.head_channels = [fold[h] * self.box_num[h] * self.box_len for h in range(3)]
        self.chan_base = [0, self.head_channels[0], self.head_channels[0] + self.head_channels[1]]
        self.ctot = sum(self.head_channels)
        self.anchors_per_image = sum(gh * gw * b for (gh, gw), b in zip(self.head_grid_sizes, self.box_num))
        self.eps = eps
        self.loss_args = dict(head_grid_sizes=self.head_grid_sizes, class_num=self.class_num, anchor_boxes=anchor_boxes,
                              iou_thresh=iou_thresh, loss_weights=loss_weights or SHIPPED_LOSS_WEIGHTS,
                              rectified_loss_weight=list(rectified_loss_weight), is_focal_loss=is_focal_loss,
                              focal_alpha=focal_alpha, focal_gamma=focal_gamma, is_tiou_recall=is_tiou_recall)
