"""Developer probe: time the HBM-bound sweeps alone (CUDA events, L2 flushed by the data size or an explicit flush) at
any shape:   python tools/sweep_probe.py --size 512 --classes 0 --batch 128 [--kind fused|dense|score|decode] [--reps 50]
Prints one JSON line: avg launch us, algorithmic bytes, GB/s, fraction of MEASURED_PEAKS.json hbm_gbs."""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import yolov3_tensorflow_b200 as yv3  # noqa: E402
from yolov3_tensorflow_b200 import synth as ps  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--classes", type=int, default=0)
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--m", type=int, default=16)
    ap.add_argument("--kind", default="fused")
    ap.add_argument("--reps", type=int, default=30)
    ap.add_argument("--thresh", type=float, default=0.5)
    ap.add_argument("--focal", type=int, default=1)
    ap.add_argument("--rect", type=int, default=1)
    ap.add_argument("--once", action="store_true", help="one launch after warm-up (for ncu)")
    a = ap.parse_args()
    anchors = ps.coco_anchors() if a.classes else ps.three_anchor_shipped()
    spec = ps.ProductSpec(a.size, a.size, a.classes, anchors, iou_thresh=0.8, is_focal_loss=bool(a.focal), focal_alpha=1.0,
                          focal_gamma=2.0, eps=1e-8)
    yv3.backend.set_epsilon(1e-8)
    dev = torch.device("cuda", 0)
    n, m = a.batch, a.m
    # random data straight on the device (same distribution as synth: N(0,1) clipped to [-8, 8])
    g = torch.Generator(device=dev).manual_seed(1234)
    h32 = a.size // 32
    pd = torch.randn((n, h32, h32, spec.ctot), generator=g, device=dev).clamp_(-8, 8)
    td = torch.from_numpy(ps.make_targets(n, a.classes, m, seed=4321, min_valid=1)).to(dev)
    loss = yv3.YOLOv3Loss(rectified_coord_num=(1 << 30) if a.rect else -1, **spec.loss_args)
    post = yv3.HeadPostProcessor(spec.head_grid_sizes, spec.class_num, spec.anchor_boxes)
    grad = torch.empty_like(pd)
    cand = torch.empty((n, post.capacity, 8), device=dev)
    counts = torch.empty((n, 3), dtype=torch.int32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if pd.numel() * 8 < (256 << 20) else None
    A, L = spec.anchors_per_image, spec.box_len

    def launch(ev):
        if a.kind == "fused":
            loss.loss_grad_and_filter(td, pd, post, a.thresh, grad_out=grad, filter_out=(cand, counts), events=ev)
        elif a.kind == "dense":
            loss.loss_and_grad(td, pd, grad_out=grad, events=ev)
        elif a.kind == "score":
            ev[0].record()
            post.filter(pd, a.thresh, out=(cand, counts))
            ev[1].record()
        elif a.kind == "decode":
            ev[0].record()
            launch.dec = launch.d.decode(pd)
            ev[1].record()
    launch.d = yv3.YOLOv3Decoder(spec.head_grid_sizes, spec.class_num, spec.anchor_boxes)
    alg = {"fused": n * (8 * A * L + 4 * m * 5), "dense": n * (4 * A * 5 + 4 * A * L + 4 * m * 5),
           "score": n * 4 * A * L, "decode": n * (4 * A * L + 4 * A * (L + 8))}[a.kind]
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.reps)]
    for _ in range(3):
        launch(evs[0])
    torch.cuda.synchronize()
    if a.once:
        launch(evs[0])
        torch.cuda.synchronize()
        return
    for k in range(a.reps):
        if flush is not None:
            flush.fill_(k & 255)
        launch(evs[k])
    torch.cuda.synchronize()
    loss.check_status()
    t = np.array([x.elapsed_time(y) for x, y in evs]) * 1e3
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    gbs = alg / (np.mean(t) * 1e-6) / 1e9
    print(json.dumps({"kind": a.kind, "size": a.size, "classes": a.classes, "batch": n, "us_mean": float(np.mean(t)), "us_min": float(np.min(t)),
                      "alg_bytes": alg, "gbs": gbs, "frac": gbs / peak, "fused": getattr(loss, "last_fused", None),
                      "l2_flushed": flush is not None}))


if __name__ == "__main__":
    main()
