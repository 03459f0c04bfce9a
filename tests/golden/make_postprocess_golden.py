"""Generates tests/golden/postprocess_*.npz by running the REFERENCE's own post-processor
(/root/reference/yolov3/yolov3_post_process.py, unmodified, imported with the removed NumPy alias
``np.float`` restored) on seeded synthetic decoded heads.  Only runnable in the build container
(the GPU box has no /root/reference); the outputs are committed.

    python tests/golden/make_postprocess_golden.py
"""
import copy
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
np.float = float          # alias removed in NumPy >= 1.24; used at yolov3_post_process.py:64,88,187
sys.path.insert(0, "/root/reference")
from yolov3.yolov3_post_process import YOLOv3PostProcessor as Ref  # noqa: E402

from oracle import ref_numpy as rn  # noqa: E402
from oracle.head_spec import SHIPPED_ANCHORS, HeadSpec, coco_anchors  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def decoded_image(spec, seed, scale=1.0, dup=False):
    rng = np.random.default_rng(seed)
    p = (rng.standard_normal((1, spec.head_grid_sizes[2][0], spec.head_grid_sizes[2][1], spec.ctot)) * scale)
    p = np.clip(p, -8, 8).astype(np.float32)
    if dup:
        # duplicate whole /32 rows so equal scores and identical boxes in different cells/heads occur
        p[0, 0, 1] = p[0, 0, 0]
        p[0, 1, 0] = p[0, 0, 0]
        # widen boxes so they overlap heavily
    dec = rn.decode(p, spec)
    return [(dec[h][1][0], dec[h][2][0]) for h in range(3)]


def run_case(name, spec, seed, score_thresh, nms_thresh, **kw):
    heads = decoded_image(spec, seed, **kw)
    args = [a for hd in heads for a in hd]
    if score_thresh == "at_score":
        # threshold exactly equal to one candidate's fp32 score -> strict '>' must drop it
        d = heads[0][0]
        sc = d[..., 4] * (np.max(d[..., 5:], axis=-1) if d.shape[-1] > 5 else 1.0)
        score_thresh = float(np.sort(sc.reshape(-1))[sc.size // 2])
    filt = Ref.filter_boxes(*args, score_thresh)
    nms = Ref.apply_nms(copy.deepcopy(filt), nms_thresh)
    # inner greedy result on the concatenated list (true survivors, ids made unique by us)
    cat = [np.concatenate([f, np.zeros((len(f), 1))], -1) if len(f) else np.zeros((0, 9)) for f in filt]
    allb = np.concatenate(cat, 0)
    allb[:, 8] = np.arange(len(allb))
    true_keep = np.array([b[8] for b in Ref._apply_nms(allb, nms_thresh)], dtype=np.int64) if len(allb) else np.zeros((0,), np.int64)
    resized = Ref.resize_boxes(copy.deepcopy(nms), np.array([spec.head_grid_sizes[0][1] * 8, spec.head_grid_sizes[0][0] * 8] * 2))
    save = dict(score_thresh=np.float64(score_thresh), nms_thresh=np.float64(nms_thresh), true_keep=true_keep)
    for h in range(3):
        save[f"pred{h}"] = heads[h][0]
        save[f"boxes{h}"] = heads[h][1]
        save[f"filt{h}"] = np.asarray(filt[h])
        save[f"nms{h}"] = np.asarray(nms[h]).reshape(-1, 9)
        save[f"resized{h}"] = np.asarray(resized[h]).reshape(-1, 9)
    np.savez_compressed(os.path.join(OUT, f"postprocess_{name}.npz"), **save)
    print(name, "K per head", [len(f) for f in filt], "kept(quirk)", [len(n) for n in nms], "true kept", len(true_keep))


if __name__ == "__main__":
    s_multi = HeadSpec.from_input(64, 96, 4, coco_anchors())
    s_zero = HeadSpec.from_input(96, 64, 0, SHIPPED_ANCHORS)
    s_big = HeadSpec.from_input(128, 128, 3, coco_anchors())
    run_case("multiclass_dense", s_multi, 11, 0.05, 0.45, scale=1.5)
    run_case("multiclass_sparse", s_multi, 12, 0.5, 0.45, scale=2.0)
    run_case("zeroclass", s_zero, 13, 0.3, 0.4)
    run_case("zeroclass_dense", s_zero, 14, 0.001, 0.3, scale=1.5)
    run_case("ties", s_multi, 15, 0.05, 0.3, dup=True, scale=1.5)
    run_case("all_empty", s_multi, 16, 0.9999, 0.45)
    run_case("at_score", s_multi, 17, "at_score", 0.45)
    run_case("bigger", s_big, 18, 0.02, 0.2, scale=1.5)
