"""TEST INFRASTRUCTURE ONLY (oracle). Not imported by the product package.

Restatement of the reference's NumPy post-processor (yolov3/yolov3_post_process.py):
  filter  :44-76  (score = max_c p_c * conf in fp32, strict ``> float32(thresh)``, raster order,
                   coords fp32 / int -> fp32, rows promoted to float64 by the int64 argmax column)
  nms     :79-129 (concat heads -> stable sort by score desc -> class-aware greedy, suppress iff
                   same class and IoU_f64 > thresh) and IoU :132-159 (float64, operation order
                   ``(a1 + a2) - inter``; 0 when overlap width or height <= 0)
  quirk   :81-104 (``start_index`` never advances, so the id column restarts at 0 per head and the
                   keep-set collides across heads; SURVEY F9 / rule T10)
  resize  :162-172

Parity status: **pinned** -- the reference module itself runs in the build container (with
``np.float = float``); tests/golden/make_postprocess_golden.py drives it and commits its outputs to
tests/golden/postprocess_*.npz, and tests/test_oracle_postprocess.py checks this restatement (and
the C twin oracle/nms_oracle.c) against those vectors bit for bit.
"""
import numpy as np


def filter_single_head(prediction, predict_boxes, score_thresh):
    """(H,W,B,L) fp32, (H,W,B,4) fp32 -> (k,8) [x0,y0,x1,y1,conf,prob,cls,score]  (:44-76)."""
    prediction = np.asarray(prediction)
    predict_boxes = np.asarray(predict_boxes)
    height, width, _, box_len = prediction.shape
    conf = prediction[..., 4]
    if box_len > 5:
        prob = np.max(prediction[..., 5:], axis=-1)                          # :57
        cls = np.argmax(prediction[..., 5:], axis=-1)                        # :58 int64, first max
        score = prob * conf                                                  # :59 fp32
    else:
        prob = np.ones_like(conf)                                            # :54
        cls = np.zeros_like(conf)                                            # :55 fp32 zeros
        score = conf
    thr = np.asarray(score_thresh).astype(score.dtype)                       # python float -> array dtype
    pos = np.nonzero((score > thr).reshape(-1))[0]                           # :62 raster (y,x,b)
    if pos.size == 0:
        return np.empty((0, 8), dtype=np.float64)                            # :64
    flat = predict_boxes.reshape(-1, 4)
    dt = predict_boxes.dtype.type
    cols = [flat[pos, 0] / dt(width), flat[pos, 1] / dt(height),             # :65-68 fp32 / int
            flat[pos, 2] / dt(width), flat[pos, 3] / dt(height),
            conf.reshape(-1)[pos], prob.reshape(-1)[pos], cls.reshape(-1)[pos], score.reshape(-1)[pos]]
    out_dtype = np.result_type(*[c.dtype for c in cols])                     # float64 iff int64 cls
    return np.stack([c.astype(out_dtype) for c in cols], axis=1)             # :74-75


def filter_boxes(p8, b8, p16, b16, p32, b32, score_thresh):
    return [filter_single_head(p8, b8, score_thresh), filter_single_head(p16, b16, score_thresh),
            filter_single_head(p32, b32, score_thresh)]


def pairwise_iou_f64(box, others):
    """_cal_iou (:132-145) of one box against (k,>=4) others, float64, same op order."""
    box = np.asarray(box, dtype=np.float64)
    others = np.asarray(others, dtype=np.float64)
    w = np.minimum(box[2], others[:, 2]) - np.maximum(box[0], others[:, 0])
    h = np.minimum(box[3], others[:, 3]) - np.maximum(box[1], others[:, 1])
    inter = w * h
    union = (box[2] - box[0]) * (box[3] - box[1]) + (others[:, 2] - others[:, 0]) * (others[:, 3] - others[:, 1]) - inter
    with np.errstate(divide="ignore", invalid="ignore"):
        iou = inter / union
    return np.where(np.logical_or(w <= 0, h <= 0), 0.0, iou)


def greedy_nms(cands, nms_thresh):
    """True greedy survivors of _apply_nms (:108-129).  cands (K,>=8) float64 rows in pre-sort order.
    Returns indices into ``cands`` in sorted (score desc, stable) order."""
    cands = np.asarray(cands, dtype=np.float64)
    k = cands.shape[0]
    order = np.argsort(-cands[:, 7], kind="stable")                          # sorted(reverse=True) is stable
    alive = np.ones(k, dtype=bool)
    keep = []
    for oi in range(k):
        i = order[oi]
        if not alive[i]:
            continue
        keep.append(i)
        later = order[oi + 1:]
        later = later[alive[later]]
        later = later[cands[later, 6] == cands[i, 6]]                        # :119 float == on class id
        if later.size:
            iou = pairwise_iou_f64(cands[i], cands[later])
            alive[later[iou > nms_thresh]] = False                           # :124 strict
    return np.asarray(keep, dtype=np.int64)


def apply_nms(boxes, nms_thresh, bug_compat=True):
    """apply_nms (:79-105).  ``boxes`` = 3 arrays (k_h,8).  Returns 3 entries: (0,9) array for an empty
    head, else a list of (9,) rows in original order.  bug_compat reproduces the id-collision quirk."""
    with_id = []
    for head_boxes in boxes:
        head_boxes = np.asarray(head_boxes)
        if len(head_boxes) == 0:
            with_id.append(np.reshape(head_boxes, (0, 9)))
            continue
        ids = np.arange(len(head_boxes), dtype=np.float64)[:, None]          # start_index stays 0 (:81-88)
        with_id.append(np.concatenate([head_boxes, ids], axis=-1))
    allb = np.concatenate(with_id, axis=0)
    keep = greedy_nms(allb, nms_thresh)
    offsets = np.cumsum([0] + [len(b) for b in with_id])
    out = []
    if bug_compat:
        kept_ids = set(allb[keep, -1].tolist())                              # :94-96
        for b in with_id:
            out.append(b if len(b) == 0 else [row for row in b if row[-1] in kept_ids])
    else:
        kept = set(keep.tolist())
        for h, b in enumerate(with_id):
            out.append(b if len(b) == 0 else [row for i, row in enumerate(b) if offsets[h] + i in kept])
    return out


def resize_boxes(boxes, target_size):
    """resize_boxes (:162-172)."""
    return [hb if len(hb) == 0 else [np.concatenate([b[:4] * target_size, b[4:]], axis=-1) for b in hb]
            for hb in boxes]
