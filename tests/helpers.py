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
    "coco_608_c80": (608, 608, 80, coco_anchors(), dict(iou_thresh=0.8)),
    # BASELINE configs[0] as named: the demo labels (classes 0..12) at 416x416
    "demo_416_c13": (416, 416, 13, coco_anchors(), dict(iou_thresh=0.8)),
    # anchors that are powers of two in grid units: with all-zero box logits every decoded box, area and intersection
    # is exact in fp32, so IoUs are the same correctly rounded quotients on any platform; iou_thresh = fl32(8 / 12)
    "pow2_128_c2": (128, 128, 2, [[(0.25, 0.25), (0.125, 0.25), (0.25, 0.125)], [(0.5, 0.5), (0.25, 0.5), (0.5, 0.25)],
                                  [(1.0, 1.0), (0.5, 1.0), (1.0, 0.5)]],
                    dict(iou_thresh=float(np.float32(8.0) / np.float32(12.0)))),
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


# ---- fixtures written by tests/golden/make_loss_golden.py (outputs of the reference's own loss / decoder / label decoder) ----
import glob as _glob
import hashlib as _hashlib
import json as _json
import os as _os

GOLDEN_DIR = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "golden")


def loss_fixture_names():
    return sorted(_os.path.basename(p)[5:-4] for p in _glob.glob(_os.path.join(GOLDEN_DIR, "loss_*.npz")))


class LossFixture:
    """one tests/golden/loss_<name>.npz: inputs + what the reference computed from them."""

    def __init__(self, name):
        self.z = np.load(_os.path.join(GOLDEN_DIR, "loss_%s.npz" % name))
        self.meta = _json.loads(str(self.z["meta"]))
        self.name = name
        self.n, self.m = self.meta["n"], self.meta["m"]
        self.rects = [bool(r) for r in self.meta["rects"]]
        H, W, C, anchors, kw = CASES[self.meta["spec_name"]]
        kw = dict(kw)
        kw.update(self.meta.get("spec_overrides", {}))
        kw.setdefault("loss_weights", SHIPPED_LOSS_WEIGHTS)
        self.spec = HeadSpec.from_input(H, W, C, anchors, eps=self.meta["eps"], **kw)
        self.targets = self.z["targets"]
        if "predicts" in self.z.files:
            self.predicts = self.z["predicts"]
        else:   # large inputs are regenerated from the seed; the hash guards against a changed NumPy stream
            self.predicts = make_predicts(self.spec, self.n, seed=self.meta["seeds"]["predicts"])
        digest = _hashlib.sha256(np.ascontiguousarray(self.predicts).tobytes()).hexdigest()
        assert digest == str(self.z["predicts_sha256"]), "regenerated inputs differ from the ones the fixture was made from"

    def __getitem__(self, key):
        return self.z[key]

    def grad(self, rect, which="64"):
        """(dense gradient in the packed shape, mask of the entries the fixture pins, complete?).  complete: every
        non-zero entry of the reference's gradient is listed, so everything outside the mask is exactly 0."""
        sfx = "_rect%d" % int(rect)
        g = np.zeros(self.predicts.size, np.float64 if which == "64" else np.float32)
        idx = self.z["grad_idx" + sfx].astype(np.int64)
        g[idx] = self.z["grad%s%s" % (which, sfx)]
        mask = np.zeros(self.predicts.size, bool)
        mask[idx] = True
        return g.reshape(self.predicts.shape), mask.reshape(self.predicts.shape), self.meta["grad_complete" + sfx]

    def decisions(self):
        """per image: valid_index and per head rows / cols / best / rho / own, as the reference decided."""
        z, out = self.z, []
        offs = z["dec_offsets"]
        for i in range(self.n):
            s, e = offs[i], offs[i + 1]
            out.append(dict(valid_index=z["dec_valid_index"][s:e],
                            heads=[dict(rows=z["dec_rows"][h, s:e], cols=z["dec_cols"][h, s:e], best=z["dec_best"][h, s:e],
                                        rho=z["dec_rho"][h, s:e], rho64=z["dec_rho64"][h, s:e], own=z["dec_own"][h, s:e],
                                        margin_best=z["dec_margin_best"][h, s:e])
                                   for h in range(3)]))
            r = z["dec_rho64"][:, s:e]
            mo = np.full(e - s, np.inf)
            for a in range(3):
                for b in range(a + 1, 3):
                    mo = np.minimum(mo, np.abs(r[a] - r[b]))
            out[-1]["margin_own"] = mo
        return out

    def mask(self, kind, h):
        """object ('obj') / background ('bg') mask of head h, (N, H, W, B) bool."""
        H, W = self.spec.head_grid_sizes[h]
        cnt = self.n * H * W * self.spec.box_num[h]
        return np.unpackbits(self.z["%s_bits%d" % (kind, h)])[:cnt].astype(bool).reshape(self.n, H, W, self.spec.box_num[h])
