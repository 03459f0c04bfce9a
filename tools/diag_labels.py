import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import make_predicts, make_spec, product_loss_args
from oracle import ref_torch as rt, ref_labels as rl
import yolov3_tensorflow_b200 as yv3
z = np.load(os.path.join(ROOT, "tests/golden/labels_test_sample.npz"))
labels = [z["flat"][5 * z["offsets"][i]:5 * z["offsets"][i + 1]] for i in range(20)]
spec = make_spec("shipped_384x480_c0")
yv3.backend.set_epsilon(spec.eps)
loss = yv3.YOLOv3Loss(rectified_coord_num=-1, **product_loss_args(spec))
L = spec.box_len
for b0 in range(0, 20, 4):
    t = rl.batch(labels[b0:b0 + 4], z["shapes"][b0:b0 + 4], (384, 480))
    p = make_predicts(spec, 4, seed=100 + b0)
    total, grad = loss.loss_and_grad(torch.from_numpy(t).cuda(), torch.from_numpy(p).cuda())
    t64, d64, g64 = rt.loss_and_grad(t, p, spec, dtype=torch.float64)
    _, d32, g32 = rt.loss_and_grad(t, p, spec, dtype=torch.float32)
    g = grad.cpu().numpy().reshape(4, -1); r = g64.reshape(4, -1); r32 = g32.reshape(4, -1)
    print("batch", b0, "detail maxdiff", np.abs(loss.loss_detail.cpu().numpy() - d64).max(), "g err", np.abs(g - r).max(),
          "g32 err", np.abs(r32 - r).max(), "gmax", np.abs(r).max(), "zero-pattern equal", np.array_equal(g == 0, r == 0))
    bad = np.argwhere(np.abs(g - r) > 1e-5 * np.abs(r) + 4 * np.abs(r32 - r).max() + 1e-6 * np.abs(r).max())
    for (i, k) in bad[:8]:
        print("   img", i, "rec", k // L, "ch", k % L, "gpu", g[i, k], "ref64", r[i, k], "ref32", r32[i, k])
    zp = np.argwhere((g == 0) != (r == 0))
    for (i, k) in zp[:8]:
        print("   zero-pattern img", i, "rec", k // L, "ch", k % L, "gpu", g[i, k], "ref64", r[i, k], "ref32", r32[i, k])
