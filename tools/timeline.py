"""Developer helper: where the two halves of consecutive steps sit in time when they run on two streams (CUDA events
at the call boundaries, eager launches, streams joined once at the end).

    python tools/timeline.py [workload] [steps] [priority: none|post|loss]
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
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
    prio = sys.argv[3] if len(sys.argv) > 3 else "post"
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
    sa = torch.cuda.Stream(priority=-1 if prio == "loss" else 0)
    sb = torch.cuda.Stream(priority=-1 if prio == "post" else 0)

    def ev():
        return torch.cuda.Event(enable_timing=True)

    for rep in range(2):
        base = ev()
        rows = []
        torch.cuda.synchronize()
        base.record()
        sa.wait_stream(torch.cuda.current_stream())
        sb.wait_stream(torch.cuda.current_stream())
        for k in range(steps):
            with torch.cuda.stream(sb):
                b0, b1, b2 = ev(), ev(), ev()
                b0.record()
                post.filter(pd, w["score_thresh"], out=(cand, counts))
                b1.record()
                post.nms(cand, counts, w["nms_thresh"], out=keep)
                b2.record()
            with torch.cuda.stream(sa):
                a0, a1 = ev(), ev()
                a0.record()
                loss.loss_and_grad(td, pd, grad_out=grad)
                a1.record()
            rows.append((a0, a1, b0, b1, b2))
        torch.cuda.synchronize()
        if rep == 0:
            continue
        print("step | loss half (assign, dense, fold)  | filter (score, gather)   | sort + nms        [us from start]")
        for k, r in enumerate(rows):
            t = [base.elapsed_time(e) * 1e3 for e in r]
            print("%4d | %7.1f .. %7.1f (%5.1f)        | %7.1f .. %7.1f (%5.1f) | .. %7.1f (%5.1f)" % (
                k, t[0], t[1], t[1] - t[0], t[2], t[3], t[3] - t[2], t[4], t[4] - t[3]))


if __name__ == "__main__":
    main()
