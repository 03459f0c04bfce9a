"""Developer helper: N plain steps of one workload (assign, dense pass, score, gather, sort, NMS x2 = 7 launches per step),
one stream, no graph -- the target of `ncu --launch-skip 14 --launch-count 7` captures.

    python tools/profile_step.py [workload] [steps]
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
import yolov3_tensorflow_b200 as yv3  # noqa: E402
from helpers import product_loss_args  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "cfg2_416_c80_b64"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    w = bench.WORKLOADS[name]
    spec = bench.build_spec(w)
    yv3.backend.set_epsilon(spec.eps)
    n, m = w["batch"], w["m"]
    p_np, t_np = bench.synth(spec, n, m, 0)
    pd, td = torch.from_numpy(p_np).cuda(), torch.from_numpy(t_np).cuda()
    loss = yv3.YOLOv3Loss(rectified_coord_num=w["rect"], **product_loss_args(spec))
    post = yv3.HeadPostProcessor(spec.head_grid_sizes, spec.class_num, spec.anchor_boxes)
    grad = torch.empty_like(pd)
    cand = torch.empty((n, post.capacity, 8), dtype=torch.float32, device="cuda")
    counts = torch.empty((n, 3), dtype=torch.int32, device="cuda")
    keep = torch.empty((n, post.capacity), dtype=torch.uint8, device="cuda")
    fused = os.environ.get("YV3_PROFILE_FUSED", "0") == "1"     # assign, fused sweep, fold, gather, sort, NMS x2
    for _ in range(steps):
        if fused:
            total, _, _, _ = loss.loss_grad_and_filter(td, pd, post, w["score_thresh"], grad_out=grad, filter_out=(cand, counts))
        else:
            total, _ = loss.loss_and_grad(td, pd, grad_out=grad)
            post.filter(pd, w["score_thresh"], out=(cand, counts))
        post.nms(cand, counts, w["nms_thresh"], out=keep)
    torch.cuda.synchronize()
    loss.check_status()
    post.check_status()
    print("ok", name, float(total), int(counts.sum()))


if __name__ == "__main__":
    main()
