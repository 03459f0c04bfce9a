"""Developer helper: one fused loss call on a named case (for compute-sanitizer / ncu)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import make_predicts, make_spec, make_targets, product_loss_args  # noqa: E402
import yolov3_tensorflow_b200 as yv3  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "tiny_32_c2"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 3
m = int(sys.argv[3]) if len(sys.argv) > 3 else 8
spec = make_spec(name)
yv3.backend.set_epsilon(spec.eps)
if os.environ.get("GOLDEN_ALL"):
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import gpu_check
    gpu_check.run_golden()
    torch.cuda.synchronize()
    print(torch.cuda.memory_summary(abbreviated=True)[:1500])
p = torch.from_numpy(make_predicts(spec, n, seed=11)).cuda()
t = torch.from_numpy(make_targets(spec, n, m, seed=12, min_valid=0, pad_in_middle=True)).cuda()
rect_num = int(sys.argv[4]) if len(sys.argv) > 4 else 0
loss = yv3.YOLOv3Loss(rectified_coord_num=rect_num, **product_loss_args(spec))
if len(sys.argv) > 5 and sys.argv[5] == "assign":
    table, iou = loss.assign(t, p)
    torch.cuda.synchronize()
    print("assign ok", int(table.sum().item()))
gold = os.environ.get("GOLDEN", "")
if gold:
    z = np.load(os.path.join(ROOT, "tests", "golden", "postprocess_multiclass_dense.npz"))
    args = []
    for h in range(3):
        args += [z["pred%d" % h], z["boxes%d" % h]]
    if "filter" in gold:
        yv3.YOLOv3PostProcessor.filter_boxes(*args, float(z["score_thresh"]))
    if "nms" in gold:
        yv3.YOLOv3PostProcessor.true_keep_mask([z["filt%d" % h].copy() for h in range(3)], float(z["nms_thresh"]))
    if "empty" in gold:
        torch.cuda.empty_cache()
    torch.cuda.synchronize()
    print("golden part done:", gold)
if os.environ.get("POISON"):
    # recycle poisoned blocks: allocate, fill with 0xFF, free -> the next torch.empty of that size reuses them
    size = loss._workspace(n, m, p.device).numel()
    loss._ws.clear()
    junk = [torch.full((size,), 255, dtype=torch.uint8, device="cuda") for _ in range(4)]
    junk2 = [torch.full((p.numel(),), float("nan"), device="cuda") for _ in range(4)]
    torch.cuda.synchronize()
    del junk, junk2
print("ptrs p %x t %x" % (p.data_ptr(), t.data_ptr()))
total, grad = loss.loss_and_grad(t, p)
print("ptrs grad %x ws %x" % (grad.data_ptr(), loss._last_ws.data_ptr()))
torch.cuda.synchronize()
print(name, "loss", total.item(), "grad abs sum", grad.abs().sum().item())
