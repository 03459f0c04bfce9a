"""Developer diagnostic (GPU box): every kernel against the oracle with error statistics printed, not
asserted.  tests/ holds the asserting versions; this script is for reading numbers after a change.

    python tools/gpu_check.py [case ...] > gpurun_out/check.log
"""
import os
import sys
import time
import traceback

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from helpers import CASES, expected_table, make_predicts, make_spec, make_targets, product_loss_args  # noqa: E402
from oracle import nms_c, ref_numpy as rn, ref_postprocess as rp, ref_torch as rt  # noqa: E402

import yolov3_tensorflow_b200 as yv3  # noqa: E402


def stats(name, got, ref, ref64=None):
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    err = np.abs(got - ref)
    denom = np.maximum(np.abs(ref), 1e-30)
    rel = err / denom
    nz = np.abs(ref) > 0
    line = "%-28s max_abs %.3e  max_rel(nz) %.3e  frac_rel>1e-5 %.2e  max|ref| %.3e  nan(got) %d" % (
        name, err.max() if err.size else 0.0, rel[nz].max() if nz.any() else 0.0,
        float((rel[nz] > 1e-5).mean()) if nz.any() else 0.0, np.abs(ref).max() if ref.size else 0.0,
        int(np.isnan(got).sum()))
    if ref64 is not None:
        e64 = np.abs(got - ref64)
        o64 = np.abs(ref - ref64)
        line += "  | vs f64: kernel %.3e oracle32 %.3e" % (e64.max(), o64.max())
    print(line, flush=True)


def section(title):
    print("\n=== " + title, flush=True)


def run_case(name, n=3, m=8):
    spec = make_spec(name)
    yv3.backend.set_epsilon(spec.eps)
    p = make_predicts(spec, n, seed=11)
    t = make_targets(spec, n, m, seed=12, min_valid=0, pad_in_middle=True)
    pd = torch.from_numpy(p).cuda()
    td = torch.from_numpy(t).cuda()
    section("%s  N=%d M=%d A=%d L=%d" % (name, n, m, spec.anchors_per_image, spec.box_len))
    skip = os.environ.get("GC_SKIP", "")
    try:
        if "decode" in skip:
            raise KeyError("skipped")
        dec = yv3.YOLOv3Decoder(spec.head_grid_sizes, spec.class_num, spec.anchor_boxes).decode(pd)
        ref = rn.decode(p, spec)
        for h in range(3):
            stats("decode raw h%d" % h, dec[h][0].cpu().numpy(), ref[h][0])
            stats("decode dec h%d" % h, dec[h][1].cpu().numpy(), ref[h][1])
            stats("decode box h%d" % h, dec[h][2].cpu().numpy(), ref[h][2])
    except Exception:
        traceback.print_exc()
    try:
        if "label" in skip:
            raise KeyError("skipped")
        lab = yv3.LabelDecoder(spec.head_grid_sizes).decode(td)
        refl = rn.label_decode(t, spec)
        print("label decode exact:", all(np.array_equal(lab[h][k].cpu().numpy(), refl[h][k]) for h in range(3) for k in range(2)))
    except Exception:
        traceback.print_exc()
    for rect in (False, True):
        try:
            loss = yv3.YOLOv3Loss(rectified_coord_num=(0 if rect else -1), **product_loss_args(spec))
            table, iou = loss.assign(td, pd)
            _, detail_ref, decisions = rn.loss(t, p, spec, rectified_active=rect)
            et, ei = expected_table(decisions, n, m)
            tb = table.cpu().numpy()
            print("rect=%s assign: table mismatches %d of %d; iou max abs diff %.3e" % (
                rect, int((tb != et).sum()), et.size, np.abs(iou.cpu().numpy() - ei).max()))
            total, grad = loss.loss_and_grad(td, pd)
            torch.cuda.synchronize()
            loss.check_status()
            total64, detail64, grad64 = rt.loss_and_grad(t, p, spec, rectified_active=rect, dtype=torch.float64)
            _, _, grad32 = rt.loss_and_grad(t, p, spec, rectified_active=rect, dtype=torch.float32)
            stats("rect=%s detail" % rect, loss.loss_detail.cpu().numpy(), detail_ref, detail64)
            print("   total kernel %.9g  oracle32 %.9g  f64 %.12g" % (total.item(), float(np.sum(detail_ref)), float(total64)))
            g = grad.cpu().numpy()
            stats("rect=%s grad" % rect, g, grad32, grad64)
            bad = np.abs(g - grad64) > 1e-5 * np.abs(grad64) + 1e-6 * np.abs(grad64).max()
            print("   grad elements outside rtol1e-5+atol1e-6*max: %d of %d; exact zeros agree: %s" % (
                int(bad.sum()), g.size, bool(np.array_equal(g == 0, grad64 == 0))))
            if bad.any():
                idx = np.argwhere(bad)[:5]
                for ix in idx:
                    ix = tuple(ix)
                    print("     at", ix, "kernel", g[ix], "o32", grad32[ix], "f64", grad64[ix])
        except Exception:
            traceback.print_exc()
    try:
        post = yv3.HeadPostProcessor(spec.head_grid_sizes, spec.class_num, spec.anchor_boxes)
        for thr, nthr in ((0.3, 0.45), (0.001, 0.45)):
            res = post.run(pd, thr, nthr)
            torch.cuda.synchronize()
            post.check_status()
            refd = rn.decode(p, spec)
            for i in range(n):
                filt = rp.filter_boxes(refd[0][1][i], refd[0][2][i], refd[1][1][i], refd[1][2][i], refd[2][1][i],
                                       refd[2][2][i], thr)
                counts = res["counts"][i].cpu().numpy()
                k = int(counts.sum())
                rows = res["cand"][i, :k].cpu().numpy().astype(np.float64)
                keep = res["keep"][i, :k].cpu().numpy().astype(bool)
                ref_rows = np.concatenate([f.reshape(-1, 8) for f in filt], 0).astype(np.float64)
                same_counts = [len(f) for f in filt] == counts.tolist()
                true_keep = np.sort(nms_c.greedy_nms(rows, nthr)) if k else np.zeros((0,), np.int64)
                ok_keep = np.array_equal(np.nonzero(keep)[0], true_keep)
                msg = "post thr=%g img%d: counts %s ref %s same=%s keep %d exact_vs_oracle_on_same_rows=%s" % (
                    thr, i, counts.tolist(), [len(f) for f in filt], same_counts, int(keep.sum()), ok_keep)
                if same_counts and k:
                    msg += " rows max_abs %.2e cls_equal %s" % (np.abs(rows - ref_rows).max(),
                                                               bool(np.array_equal(rows[:, 6], ref_rows[:, 6])))
                print(msg, flush=True)
    except Exception:
        traceback.print_exc()


def run_golden():
    import glob
    section("post-process golden fixtures (reference's own outputs)")
    for path in sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "postprocess_*.npz"))):
        z = np.load(path)
        try:
            args = []
            for h in range(3):
                args += [z["pred%d" % h], z["boxes%d" % h]]
            filt = yv3.YOLOv3PostProcessor.filter_boxes(*args, float(z["score_thresh"]))
            okf = all(filt[h].shape == z["filt%d" % h].shape and filt[h].dtype == z["filt%d" % h].dtype and
                      np.array_equal(filt[h], z["filt%d" % h]) for h in range(3))
            ref_filt = [z["filt%d" % h].copy() for h in range(3)]
            mask = yv3.YOLOv3PostProcessor.true_keep_mask(ref_filt, float(z["nms_thresh"]))
            okk = np.array_equal(np.nonzero(mask)[0], np.sort(z["true_keep"]))
            nms = yv3.YOLOv3PostProcessor.apply_nms(ref_filt, float(z["nms_thresh"]))
            okn = all(np.array_equal(np.asarray(nms[h]).reshape(-1, 9), z["nms%d" % h]) for h in range(3))
            print("%-32s filter_exact=%s true_keep_exact=%s apply_nms_exact=%s" % (os.path.basename(path), okf, okk, okn),
                  flush=True)
        except Exception:
            traceback.print_exc()


if __name__ == "__main__":
    names = sys.argv[1:] or ["tiny_32_c2", "coco_64x96_c4", "coco_128_c3_focal_tiou", "shipped_384x480_c0",
                             "single_160_c1_focal"]
    print(torch.cuda.get_device_name(0))
    t0 = time.time()
    skip = os.environ.get("GC_SKIP", "")
    if "golden" not in skip:
        run_golden()
    for nm in names:
        run_case(nm)
    print("done in %.1fs" % (time.time() - t0))
