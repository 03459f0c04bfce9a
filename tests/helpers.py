"""Shared builders for the parity tests: one description -> (oracle HeadSpec, product constructor args)."""
import numpy as np

from oracle.head_spec import SHIPPED_ANCHORS, SHIPPED_LOSS_WEIGHTS, HeadSpec, coco_anchors


def three_anchor_shipped():
    a = [list(h) for h in SHIPPED_ANCHORS]
    a[1] = a[1] + [SHIPPED_ANCHORS[2][1]]
    return a


CASES = {
    # name: (H, W, class_num, anchors, kwargs)
    "shipped_384x480_c0": (384, 480, 0, SHIPPED_ANCHORS, dict(iou_thresh=0.8)),
    "coco_64x96_c4": (64, 96, 4, coco_anchors(), dict(iou_thresh=0.5)),
    "coco_128_c3_focal_tiou": (128, 128, 3, coco_anchors(),
                               dict(iou_thresh=0.6, is_focal_loss=True, focal_alpha=0.7, focal_gamma=2.0,
                                    is_tiou_recall=True)),
    "tiny_32_c2": (32, 32, 2, SHIPPED_ANCHORS, dict(iou_thresh=0.3)),
    # record lengths >= 12 floats take the whole-sector gradient path of the dense pass: even and odd L
    "coco_96x64_c13": (96, 64, 13, coco_anchors(), dict(iou_thresh=0.5)),
    "coco_64_c20": (64, 64, 20, coco_anchors(), dict(iou_thresh=0.4)),
    "coco_416_c80": (416, 416, 80, coco_anchors(), dict(iou_thresh=0.8)),
    "single_512_c0_focal": (512, 512, 0, three_anchor_shipped(),
                            dict(iou_thresh=0.8, is_focal_loss=True, focal_alpha=1.0, focal_gamma=2.0,
                                 rectified_loss_weight=(1.0, 1.0, 1.0))),
    "single_160_c1_focal": (160, 160, 1, three_anchor_shipped(),
                            dict(iou_thresh=0.7, is_focal_loss=True, focal_alpha=1.0, focal_gamma=2.0,
                                 rectified_loss_weight=(1.0, 1.0, 1.0))),
}


def make_spec(name, eps=1e-8):
    H, W, C, anchors, kw = CASES[name]
    kw = dict(kw)
    kw.setdefault("loss_weights", SHIPPED_LOSS_WEIGHTS)
    return HeadSpec.from_input(H, W, C, anchors, eps=eps, **kw)


def product_loss_args(spec):
    """positional/keyword arguments of YOLOv3Loss.__init__ for a HeadSpec."""
    return dict(head_grid_sizes=spec.head_grid_sizes, class_num=spec.class_num, anchor_boxes=spec.anchor_boxes,
                iou_thresh=spec.iou_thresh, loss_weights=spec.loss_weights,
                rectified_loss_weight=list(spec.rectified_loss_weight), is_focal_loss=spec.is_focal_loss,
                focal_alpha=spec.focal_alpha, focal_gamma=spec.focal_gamma, is_tiou_recall=spec.is_tiou_recall)


def make_predicts(spec, n, seed, scale=1.0, clip=8.0):
    rng = np.random.default_rng(seed)
    h32, w32 = spec.head_grid_sizes[2]
    p = rng.standard_normal((n, h32, w32, spec.ctot)).astype(np.float32) * np.float32(scale)
    return np.clip(p, -clip, clip).astype(np.float32)


def make_targets(spec, n, m, seed, min_valid=0, max_valid=None, pad_in_middle=False):
    rng = np.random.default_rng(seed)
    max_valid = m if max_valid is None else max_valid
    t = np.full((n, m, 5), -1.0, dtype=np.float32)
    for i in range(n):
        v = int(rng.integers(min_valid, max_valid + 1))
        if v == 0:
            continue
        xy = rng.uniform(0.02, 0.98, size=(v, 2))
        wh = np.exp(rng.uniform(np.log(0.02), np.log(0.6), size=(v, 2)))
        wh = np.minimum(wh, 2.0 * np.minimum(xy, 1.0 - xy))
        cls = rng.integers(0, max(spec.class_num, 1), size=(v, 1))
        rows = np.concatenate([xy, wh, cls], axis=1).astype(np.float32)
        slots = np.sort(rng.choice(m, size=v, replace=False)) if pad_in_middle else np.arange(v)
        t[i, slots] = rows
    return t.reshape(n, m * 5)


def expected_table(decisions, n, m):
    """oracle decisions -> the (N,M,3,4) int32 table and (N,M,3) IoU of yv3_assign."""
    table = np.zeros((n, m, 3, 4), np.int32)
    table[..., 0:3] = -1
    iou = np.zeros((n, m, 3), np.float32)
    for i in range(n):
        d = decisions[i]
        vidx = d["valid_index"]
        for h in range(3):
            hd = d["heads"][h]
            table[i, vidx, h, 0] = hd["rows"]
            table[i, vidx, h, 1] = hd["cols"]
            table[i, vidx, h, 2] = hd["best"]
            table[i, vidx, h, 3] = hd["own"].astype(np.int32)
            iou[i, vidx, h] = hd["rho"]
    return table, iou
