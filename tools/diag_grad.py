"""Developer helper: where does the fused gradient differ from the fp64 oracle?  (GPU box; tests/ only infrastructure)"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import make_predicts, make_spec, make_targets, product_loss_args
from oracle import ref_torch as rt
import yolov3_tensorflow_b200 as yv3
name = sys.argv[1] if len(sys.argv) > 1 else "coco_416_c80"
n, m = 3, 8
spec = make_spec(name)
yv3.backend.set_epsilon(spec.eps)
p = make_predicts(spec, n, seed=11)
t = make_targets(spec, n, m, seed=12, min_valid=0, pad_in_middle=True).reshape(n, m, 5)
t[1] = -1.0
t[2, 1] = t[2, 0]
t = t.reshape(n, m * 5)
loss = yv3.YOLOv3Loss(rectified_coord_num=-1, **product_loss_args(spec))
total, grad = loss.loss_and_grad(torch.from_numpy(t).cuda(), torch.from_numpy(p).cuda())
total64, detail64, grad64 = rt.loss_and_grad(t, p, spec, dtype=torch.float64)
g = grad.cpu().numpy().reshape(n, -1)
r = grad64.reshape(n, -1)
L = spec.box_len
bad = np.abs(g - r) > 1e-5 * np.abs(r) + 1e-6 * np.abs(r).max()
print("bad", bad.sum(), "of", bad.size, "gmax", np.abs(r).max())
idx = np.argwhere(bad)
for (i, k) in idx[:40]:
    print("img", i, "rec", k // L, "ch", k % L, "gpu", g[i, k], "ref", r[i, k])
