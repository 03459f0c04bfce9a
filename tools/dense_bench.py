"""Developer helper: per-stage device times of one workload (CUDA events), plus reference points for the roofline:
a cudaMemset of the gradient tensor (pure-write ceiling) and a device copy of it (read+write ceiling).

    python tools/dense_bench.py [workload] [steps] [--l2gran 32|64|128]
"""
import ctypes
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
import yolov3_tensorflow_b200 as yv3  # noqa: E402
from helpers import product_loss_args  # noqa: E402


def timed(fn, steps, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for a, b in ev:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    t = sorted(a.elapsed_time(b) for a, b in ev)
    return {"median_us": 1e3 * t[len(t) // 2], "p10_us": 1e3 * t[len(t) // 10], "p90_us": 1e3 * t[(9 * len(t)) // 10]}


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    name = args[0] if args else "cfg2_416_c80_b64"
    steps = int(args[1]) if len(args) > 1 else 100
    if "--l2gran" in sys.argv:
        gran = int(sys.argv[sys.argv.index("--l2gran") + 1])
        rt = ctypes.CDLL("libcudart.so.12")
        torch.cuda.init()
        torch.zeros(1, device="cuda")
        rc = rt.cudaDeviceSetLimit(5, ctypes.c_size_t(gran))          # cudaLimitMaxL2FetchGranularity
        v = ctypes.c_size_t(0)
        rt.cudaDeviceGetLimit(ctypes.byref(v), 5)
        print("cudaLimitMaxL2FetchGranularity rc", rc, "now", v.value)
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
    out = {"workload": name}
    # dense sweep alone: events inside loss_and_grad
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for _ in range(5):
        loss.loss_and_grad(td, pd, grad_out=grad)
    torch.cuda.synchronize()
    for e in ev:
        loss.loss_and_grad(td, pd, events=e, grad_out=grad)
    torch.cuda.synchronize()
    t = sorted(a.elapsed_time(b) for a, b in ev)
    out["loss_dense"] = {"median_us": 1e3 * t[len(t) // 2], "p10_us": 1e3 * t[len(t) // 10]}
    out["loss_all"] = timed(lambda: loss.loss_and_grad(td, pd, grad_out=grad), steps)
    out["loss_fwd_only"] = timed(lambda: loss.loss(td, pd), steps)
    from oracle import ref_numpy as rn
    heads = [torch.from_numpy(np.ascontiguousarray(h.reshape(n, h.shape[1], h.shape[2], -1))).cuda() for h in rn.unpack(p_np, spec)]
    hgrads = [torch.empty_like(h) for h in heads]
    out["loss_all_unpacked_heads"] = timed(lambda: loss.loss_and_grad_heads(td, heads, grads_out=hgrads), steps)
    out["pack_copy_for_context"] = timed(lambda: torch.cat([h.reshape(n, spec.head_grid_sizes[2][0], spec.head_grid_sizes[2][1], -1) for h in heads], dim=-1), steps)
    out["filter"] = timed(lambda: post.filter(pd, w["score_thresh"], out=(cand, counts)), steps)
    out["nms"] = timed(lambda: post.nms(cand, counts, w["nms_thresh"], out=keep), steps)
    out["memset_grad"] = timed(lambda: grad.zero_(), steps)
    out["copy_grad"] = timed(lambda: grad.copy_(pd), steps)
    gb = grad.numel() * 4 / 1e9
    out["memset_GBps"] = gb / (out["memset_grad"]["median_us"] * 1e-6)
    out["copy_GBps"] = 2 * gb / (out["copy_grad"]["median_us"] * 1e-6)
    loss.check_status()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
