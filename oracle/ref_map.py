"""TEST INFRASTRUCTURE ONLY (oracle). Not imported by the product package.

AP / mAP the way the tool the reference points to computes it: run.py:78-79 and README.md:137 name Cartucho/mAP
(https://github.com/Cartucho/mAP) as the evaluation step the author left as a TODO.  The tool is a third-party
dependency that is NOT under /root/reference (no submodule, no version pinned -- a bare URL), so this file restates the
algorithm of its published ``main.py`` in plain Python floats:

  * per class, detections of all images sorted by confidence descending (``sorted(..., reverse=True)``: stable, ties keep
    file order = image order, then line order);
  * per detection: over the ground-truth boxes of the same image and class, IoU with inclusive pixel widths
    (``iw = min(x1) - max(x0) + 1``), the LARGEST wins (``ov > ovmax``: first on ties); true positive iff
    ``ovmax >= MINOVERLAP`` (0.5) and that box is not ``used`` yet, otherwise false positive;
  * ``voc_ap``: cumulative fp / tp, rec = tp / #gt, prec = tp / (fp + tp); mrec = [0] + rec + [1], mpre = [0] + prec + [0];
    mpre made monotone from the right; AP = sum over i with mrec[i] != mrec[i-1] of (mrec[i] - mrec[i-1]) * mpre[i];
  * mAP = sum of AP / number of classes that have ground truth.

Parity status: **parity unpinned** -- the tool cannot be fetched here (no network) and the reference holds no call site,
test or golden vector for it (it is a TODO there).  Ground-truth boxes come from the label batch the loss takes:
(x -+ w/2) * image_w, (y -+ h/2) * image_h in float64 (the tool reads them from text files the user writes).
"""
import numpy as np


def gt_boxes(targets_row, image_w, image_h, class_agnostic=False):
    """(M*5,) fp32 label row -> list of (class, [x0, y0, x1, y1]) in pixels for the valid labels (x >= 0)."""
    t = np.asarray(targets_row, dtype=np.float32).reshape(-1, 5)
    out = []
    for x, y, w, h, c in t:
        if not x >= 0:
            continue
        x, y, w, h = float(x), float(y), float(w), float(h)
        hw, hh = w * 0.5, h * 0.5
        out.append((0 if class_agnostic else int(c), [(x - hw) * image_w, (y - hh) * image_h, (x + hw) * image_w, (y + hh) * image_h]))
    return out


def voc_ap(rec, prec):
    mrec = [0.0] + list(rec) + [1.0]
    mpre = [0.0] + list(prec) + [0.0]
    for i in range(len(mpre) - 2, -1, -1):
        mpre[i] = max(mpre[i], mpre[i + 1])
    ap = 0.0
    for i in range(1, len(mrec)):
        if mrec[i] != mrec[i - 1]:
            ap += (mrec[i] - mrec[i - 1]) * mpre[i]
    return ap


def mean_ap(detections, ground_truth, num_classes, min_overlap=0.5, inclusive=True):
    """detections: per image a list of (class, score, [x0, y0, x1, y1]) in file order; ground_truth: per image the list of
    gt_boxes().  Returns (ap per class with -1 where a class has no ground truth, mAP, gt count per class)."""
    inc = 1.0 if inclusive else 0.0
    gt_count = [0] * num_classes
    for boxes in ground_truth:
        for c, _ in boxes:
            if 0 <= c < num_classes:
                gt_count[c] += 1
    ap = [-1.0] * num_classes
    for c in range(num_classes):
        dets = [(s, i, bb) for i, img in enumerate(detections) for (cc, s, bb) in img if cc == c]
        dets = sorted(dets, key=lambda d: d[0], reverse=True)                 # stable
        used = [[False] * len(g) for g in ground_truth]
        tp, fp = [0] * len(dets), [0] * len(dets)
        for k, (s, i, bb) in enumerate(dets):
            ovmax, match = -1.0, -1
            for j, (gc, bbgt) in enumerate(ground_truth[i]):
                if gc != c:
                    continue
                bi = [max(bb[0], bbgt[0]), max(bb[1], bbgt[1]), min(bb[2], bbgt[2]), min(bb[3], bbgt[3])]
                iw = bi[2] - bi[0] + inc
                ih = bi[3] - bi[1] + inc
                if iw > 0 and ih > 0:
                    ua = (bb[2] - bb[0] + inc) * (bb[3] - bb[1] + inc) + (bbgt[2] - bbgt[0] + inc) * (bbgt[3] - bbgt[1] + inc) - iw * ih
                    ov = iw * ih / ua
                    if ov > ovmax:
                        ovmax, match = ov, j
            if ovmax >= min_overlap and not used[i][match]:
                tp[k] = 1
                used[i][match] = True
            else:
                fp[k] = 1
        if gt_count[c] == 0:
            continue
        cum_fp, cum_tp = np.cumsum(fp).tolist(), np.cumsum(tp).tolist()
        rec = [float(t) / gt_count[c] for t in cum_tp]
        prec = [float(t) / (f + t) for t, f in zip(cum_tp, cum_fp)]
        ap[c] = voc_ap(rec, prec)
    have = [a for a in ap if a >= 0]
    m = 0.0
    for a in have:
        m += a
    return np.asarray(ap), (m / len(have) if have else 0.0), np.asarray(gt_count, np.int32)
