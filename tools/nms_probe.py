"""Developer probe: the NMS call alone (sort + NMS kernels) on one workload's candidates, timed with CUDA events.
    [YV3_NMS_SKIP=bits] python tools/nms_probe.py [workload] [reps]
YV3_NMS_SKIP bits (timing experiments only, results are then incomplete): 1 no sort kernel, 2 no warp NMS, 4 no
long-segment kernel, 8 launch the queue-driven kernels from the host instead of from the device."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
import yolov3_tensorflow_b200 as yv3  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "cfg4_416_c80_b256_dense"
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    w = bench.WORKLOADS[name]
    spec = bench.product_spec(w)
    yv3.backend.set_epsilon(spec.eps)
    n = bench.per_rank_batch(w, 1)
    p_np, _ = bench.synth(spec, n, w["m"], 0)
    pd = torch.from_numpy(p_np).cuda()
    post = yv3.HeadPostProcessor(spec.head_grid_sizes, spec.class_num, spec.anchor_boxes)
    cand, counts = post.filter(pd, w["score_thresh"])
    keep = torch.empty((n, post.capacity), dtype=torch.uint8, device="cuda")
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for _ in range(3):
        post.nms(cand, counts, w["nms_thresh"], out=keep)
    torch.cuda.synchronize()
    for a, b in evs:
        a.record()
        post.nms(cand, counts, w["nms_thresh"], out=keep)
        b.record()
    torch.cuda.synchronize()
    t = np.array([a.elapsed_time(b) for a, b in evs]) * 1e3
    print("%s skip=%s  nms call: mean %.1f us  min %.1f us  (%.0f candidates per image)" % (
        name, os.environ.get("YV3_NMS_SKIP", "0"), t.mean(), t.min(), float(counts.sum()) / n))


if __name__ == "__main__":
    main()
