"""Developer helper: one pass over EVERY kernel of the library at a small shape and at 416x416/80 -- the target of
   compute-sanitizer --tool memcheck|racecheck|synccheck|initcheck python tools/sanitize_run.py
(one tool per gpurun call; summaries go to profiles/)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import yolov3_tensorflow_b200 as yv3  # noqa: E402
from helpers import make_predicts, make_spec, make_targets, product_loss_args  # noqa: E402


def one(name, n, m, thr):
    spec = make_spec(name)
    yv3.backend.set_epsilon(spec.eps)
    p = torch.from_numpy(make_predicts(spec, n, seed=1)).cuda()
    t = torch.from_numpy(make_targets(spec, n, m, seed=2, min_valid=0, pad_in_middle=True)).cuda()
    H, W = spec.head_grid_sizes[0][0] * 8, spec.head_grid_sizes[0][1] * 8
    post = yv3.HeadPostProcessor(spec.head_grid_sizes, spec.class_num, spec.anchor_boxes)
    loss = yv3.YOLOv3Loss(rectified_coord_num=10 ** 9, **product_loss_args(spec))
    loss.loss_and_grad(t, p)                                                    # assign, dense sweep, fold
    loss.loss(t, p)                                                             # forward only
    loss.loss_grad_and_filter(t, p, post, thr)                                  # fused sweep, gather
    heads = [p[..., spec.chan_base[h]:spec.chan_base[h] + spec.head_channels[h]].reshape(
        n, spec.head_grid_sizes[h][0], spec.head_grid_sizes[h][1], -1).contiguous() for h in range(3)]
    loss.loss_and_grad_heads(t, heads)
    loss.loss_grad_and_filter_heads(t, heads, post, thr)
    post.filter_heads(heads, thr)
    res = post.run(p, thr, 0.45)                                                # score sweep, gather, sort (+ dispatch), NMS
    rows, kc = post.collect(res, (W, H))
    acc = yv3.MeanAPAccumulator(spec.class_num, (W, H), max_detections=1 << 16)
    acc.update_rows(rows, kc, t)
    acc.compute()
    yv3.YOLOv3Decoder(spec.head_grid_sizes, spec.class_num, spec.anchor_boxes).decode(p)
    yv3.LabelDecoder(spec.head_grid_sizes).decode(t)
    loss.assign(t, p)
    torch.cuda.synchronize()
    loss.check_status()
    post.check_status()
    return spec, post


def nms_paths(post):
    """segments of every length class: <= 32 (inside the sort kernel), 33..256 (warp kernel), > 256 (long-segment kernels)"""
    rng = np.random.default_rng(3)
    lens = [5, 32, 33, 200, 256, 300, 1100]
    rows = []
    for c, ln in enumerate(lens):
        cx, cy = rng.uniform(0.3, 0.7, ln), rng.uniform(0.3, 0.7, ln)
        w, h = rng.uniform(0.05, 0.3, ln), rng.uniform(0.05, 0.3, ln)
        r = np.zeros((ln, 8), np.float32)
        r[:, 0], r[:, 1], r[:, 2], r[:, 3] = cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2
        r[:, 4] = rng.uniform(0.2, 1, ln)
        r[:, 5] = 1
        r[:, 6] = c
        r[:, 7] = r[:, 4]
        rows.append(r)
    rows = np.concatenate(rows)
    k = len(rows)
    cand = np.zeros((2, post.capacity, 8), np.float32)
    cand[:, :k] = rows[rng.permutation(k)]
    counts = np.array([[k // 2, k // 4, k - k // 2 - k // 4]] * 2, np.int32)
    post.nms(torch.from_numpy(cand).cuda(), torch.from_numpy(counts).cuda(), 0.45)
    torch.cuda.synchronize()


def main():
    one("coco_64x96_c4", 3, 6, 0.05)
    one("shipped_384x480_c0", 2, 5, 0.5)            # class_num 0: spans of four tiles, single-class NMS (cluster kernel)
    spec, post = one("coco_416_c80", 2, 8, 0.05)
    nms_paths(post)
    lb = yv3.LabelBatcher((416, 416))
    lb.batch([np.array([0.5, 0.5, 0.2, 0.2, 1], np.float32), np.zeros((0,), np.float32)], [(300, 400), (416, 416)], slots=3)
    torch.cuda.synchronize()
    print("sanitize_run ok")


if __name__ == "__main__":
    main()
