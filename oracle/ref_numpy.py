"""TEST INFRASTRUCTURE ONLY (oracle). Not imported by the product package.

Op-for-op NumPy restatement (fp32 by default, fp64 on request) of the reference's
TensorFlow head path, returning every integer decision and its margin, plus the
closed-form backward that TF autodiff produces.

Follows, line by line:
  packing     yolov3/yolov3_detector.py:80-85
  decoder     yolov3/yolov3_decoder.py:48-60 (left_top), :89-117 (_unpack), :119-192
  labels      yolov3/label_decoder.py:26-60
  loss        yolov3/yolov3_loss.py:81-138 (loss), :140-164 (rectified), :166-222,
              :225-252 (_get_valid_target), :254-303 (_calc_iou), :305-369 (_single_head_loss)
  epsilon     run.py:26   padding  dataset/file_util.py:97

Parity status: **pinned** -- tests/test_oracle_pinned.py holds this file to tests/golden/loss_*.npz, the outputs of
the reference's own yolov3_loss.py / yolov3_decoder.py / label_decoder.py run unmodified under the torch-backed
``tensorflow`` stand-in (oracle/tf_standin, generator tests/golden/make_loss_golden.py): integer decisions and masks
bit for bit, fp32 values to the ulp differences of NumPy's vs torch's exp/log and summation order, fp64 <= 1e-12.
What stays assumed about TensorFlow itself is listed in oracle/tf_standin/tensorflow/__init__.py (A1-A6) and
SURVEY.md Appendix B (clip gradient inclusive, argmax first index, softmax = exp(x-max) * (1/sum), python scalars
rounded once to fp32).
"""
import numpy as np

from .head_spec import HEAD_FOLD, HeadSpec

NEG_INF = -np.inf


# --------------------------------------------------------------------------- layout
def pack_heads(h8, h16, h32):
    """(N,H,W,B*L) x3 -> (N,H32,W32,Ctot): row-major Reshape + channel concat
    (yolov3_detector.py:80-85)."""
    n, h32h, h32w = h32.shape[0], h32.shape[1], h32.shape[2]
    return np.concatenate([h8.reshape(n, h32h, h32w, -1), h16.reshape(n, h32h, h32w, -1),
                           h32.reshape(n, h32h, h32w, -1)], axis=-1)


def unpack(predicts, spec: HeadSpec):
    """(N,H32,W32,Ctot) -> [(N,H,W,B,L)]x3  (yolov3_decoder.py:99-117)."""
    out = []
    n = predicts.shape[0]
    for h in range(3):
        H, W = spec.head_grid_sizes[h]
        s, e = spec.chan_base[h], spec.chan_base[h] + spec.head_channels[h]
        out.append(predicts[..., s:e].reshape(n, H, W, spec.box_num[h], spec.box_len))
    return out


def packed_offset(spec: HeadSpec, n, h, y, x, b, ch):
    """float offset of (image n, head h, cell (y,x), anchor b, channel ch) -- SURVEY Appendix C."""
    H, W = spec.head_grid_sizes[h]
    g = y * W + x
    m = HEAD_FOLD[h]
    row = g // m
    chan = spec.chan_base[h] + (g % m) * spec.box_num[h] * spec.box_len + b * spec.box_len + ch
    return (n * spec.rows32 + row) * spec.ctot + chan


# --------------------------------------------------------------------------- decode
def _const(v, dtype):
    """a graph constant: rounded once to fp32 (convert_to_tensor(v, float32)), then widened for the fp64 twin."""
    return np.asarray(v, dtype=np.float32).astype(dtype)[()]


def _sigmoid(x):
    one = x.dtype.type(1)
    return one / (one + np.exp(-x))


def clip_bounds(spec: HeadSpec, dtype):
    """python floats eps and 1-eps rounded once to fp32 (Appendix B), widened in the fp64 twin."""
    return dtype(np.float32(spec.eps)), dtype(np.float32(1.0 - spec.eps))


def decode_head(raw, spec: HeadSpec, h, dtype=np.float32):
    """raw (N,H,W,B,L) -> decoded (N,H,W,B,L), boxes (N,H,W,B,4)  (yolov3_decoder.py:119-192)."""
    dtype = np.dtype(dtype).type
    raw = raw.astype(dtype, copy=False)
    H, W = spec.head_grid_sizes[h]
    lo, hi = clip_bounds(spec, dtype)
    anchors = spec.anchors_grid(dtype)[h]                         # (B,2)
    # left_top[y, x] = (x, y)  (:48-60)
    lt = np.stack(np.meshgrid(np.arange(W), np.arange(H)), axis=-1).reshape(H, W, 1, 2).astype(dtype)
    xy = np.maximum(np.minimum(_sigmoid(raw[..., 0:2]), hi), lo) + lt      # :153-155
    wh = np.exp(raw[..., 2:4]) * anchors                                     # :167-168
    conf = np.maximum(np.minimum(_sigmoid(raw[..., 4:5]), hi), lo)           # :178-179
    parts = [xy, wh, conf]
    if spec.class_num >= 1:                                                   # :132
        z = raw[..., 5:]
        z = z - np.max(z, axis=-1, keepdims=True)                            # :189-190
        # tf.nn.softmax (CPU functor): exp(z - max z) * (1 / sum)
        e = np.exp(z - np.max(z, axis=-1, keepdims=True))
        q = e * (dtype(1) / np.sum(e, axis=-1, keepdims=True))
        parts.append(np.maximum(np.minimum(q, hi), lo))                      # :191
    decoded = np.concatenate(parts, axis=-1)
    half = wh / dtype(2)
    boxes = np.concatenate([xy - half, xy + half], axis=-1)                  # :138-139
    return decoded, boxes


def decode(predicts, spec: HeadSpec, dtype=np.float32):
    """YOLOv3Decoder.decode: [(raw[...,0:4], decoded, boxes)] in /8,/16,/32 order (:62-87)."""
    raws = unpack(np.asarray(predicts), spec)
    out = []
    for h in range(3):
        d, b = decode_head(raws[h], spec, h, dtype)
        out.append((raws[h][..., 0:4].astype(dtype, copy=False), d, b))
    return out


def label_decode(targets, spec: HeadSpec, dtype=np.float32):
    """LabelDecoder.decode: (N,M*5) -> [(targets_h (N,M,5), boxes_h (N,M,4))]x3 (label_decoder.py:26-60)."""
    dtype = np.dtype(dtype).type
    t = np.asarray(targets).astype(dtype, copy=False)
    t = t.reshape(t.shape[0], -1, 5)
    out = []
    for h in range(3):
        H, W = spec.head_grid_sizes[h]
        wh_h = np.asarray([W, H], dtype=dtype)
        xy = t[:, :, 0:2] * wh_h
        wh = t[:, :, 2:4] * wh_h
        th = np.concatenate([xy, wh, t[:, :, 4:5]], axis=-1)
        half = wh / dtype(2)
        out.append((th, np.concatenate([xy - half, xy + half], axis=-1)))
    return out


# --------------------------------------------------------------------------- IoU / assignment
def _iou(pbox, parea, tbox, tarea, tiou):
    """broadcasting IoU exactly as yolov3_loss.py:276-293."""
    lt = np.maximum(pbox[..., 0:2], tbox[..., 0:2])
    rb = np.minimum(pbox[..., 2:4], tbox[..., 2:4])
    wh = np.maximum(rb - lt, pbox.dtype.type(0))
    inter = wh[..., 0] * wh[..., 1]
    with np.errstate(divide="ignore", invalid="ignore"):
        iou = inter / (parea + tarea - inter)
        if tiou:
            iou = iou * inter / tarea
    return iou


def calc_iou(target, tboxes, pred, pboxes, spec: HeadSpec):
    """_calc_iou (:254-303) for one image and one head.
    Returns max_iou (H,W,B), response_iou (V,B), rows, cols (V,) int32."""
    V = target.shape[0]
    parea = pred[..., 2] * pred[..., 3]                                      # :267
    grid = np.floor(target[:, 0:2]).astype(np.int32)                         # :269 (col,row)
    cols, rows = grid[:, 0], grid[:, 1]                                      # :270 reversed
    tarea = target[:, 2] * target[:, 3]                                      # :273
    if V == 0:
        max_iou = np.full(parea.shape, NEG_INF, dtype=pred.dtype)            # reduce_max over empty
        return max_iou, np.zeros((0, pred.shape[2]), pred.dtype), rows, cols
    iou = _iou(pboxes[:, :, :, None, :], parea[..., None], tboxes[None, None, None], tarea, spec.is_tiou_recall)
    max_iou = np.max(iou, axis=-1)                                           # :294
    H, W = parea.shape[0], parea.shape[1]
    if np.any(rows < 0) or np.any(rows >= H) or np.any(cols < 0) or np.any(cols >= W):
        raise IndexError("GT centre outside the grid (tf.gather_nd would raise on CPU)")
    rbox = pboxes[rows, cols]                                                # (V,B,4) :281
    rarea = parea[rows, cols]                                                # (V,B)   :271
    riou = _iou(rbox, rarea, tboxes[:, None, :], tarea[:, None], spec.is_tiou_recall)   # :296-299
    return max_iou, riou, rows, cols


def _margin_sorted(values, axis=-1):
    """gap between the largest and second largest entry (inf if only one)."""
    if values.shape[axis] < 2:
        return np.full(values.shape[:axis] + values.shape[axis:][1:], np.inf)
    s = np.sort(values.astype(np.float64), axis=axis)
    return np.take(s, -1, axis=axis) - np.take(s, -2, axis=axis)


def single_image(dec, lab, spec: HeadSpec, dtype):
    """_single_image_loss (:166-222). dec[h]=(decoded(H,W,B,L), boxes), lab[h]=(target(M,5), tboxes(M,4)).
    Returns loss (5,3) and a dict of decisions/margins."""
    valid = lab[0][0][:, 0] >= 0                                             # :239 (head /8 x)
    vidx = np.nonzero(valid)[0]                                              # :241 order preserved
    V = len(vidx)
    per_head = []
    for h in range(3):
        tgt, tb = lab[h][0][vidx], lab[h][1][vidx]
        max_iou, riou, rows, cols = calc_iou(tgt, tb, dec[h][0], dec[h][1], spec)
        if V:
            rho = np.max(riou, axis=-1)                                       # :300
            best = np.argmax(riou, axis=-1).astype(np.int32)                 # :301 first index
        else:
            rho = np.zeros((0,), dtype)
            best = np.zeros((0,), np.int32)
        per_head.append(dict(tgt=tgt, max_iou=max_iou, riou=riou, rows=rows, cols=cols, rho=rho, best=best))
    rhos = [p["rho"] for p in per_head]
    own = [np.logical_and(rhos[h] >= rhos[(h + 1) % 3], rhos[h] >= rhos[(h + 2) % 3]) for h in range(3)]  # :203-208

    loss = np.zeros((5, 3), dtype)
    decisions = dict(valid_index=vidx.astype(np.int32), heads=[])
    lw = _const(spec.loss_weights, dtype)                                    # (3,5): xy, wh, noobj, obj, cls
    thresh = _const(spec.iou_thresh, dtype)
    gamma, alpha = _const(spec.focal_gamma, dtype), _const(spec.focal_alpha, dtype)
    for h in range(3):
        p = per_head[h]
        H, W = spec.head_grid_sizes[h]
        B = spec.box_num[h]
        pos = np.nonzero(own[h])[0]                                          # tf.where: ascending
        rows, cols, best = p["rows"][pos], p["cols"][pos], p["best"][pos]
        obj = np.zeros((H, W, B), bool)
        obj[rows, cols, best] = True                                         # :328 sparse_to_dense
        bg = np.logical_and(p["max_iou"] < thresh, ~obj)                     # :331-332
        decoded = dec[h][0]
        conf = decoded[..., 4]
        one = dtype(1)
        with np.errstate(divide="ignore", invalid="ignore"):
            noobj = -np.log(one - conf)                                      # :335
            if spec.is_focal_loss:
                noobj = noobj * np.power(conf, gamma)      # :337
            noobj_l = lw[h, 2] * np.sum(noobj * bg.astype(dtype))            # :338
            rt = p["tgt"][pos]                                               # :341
            rp = decoded[rows, cols, best]                                   # :342 (R,L)
            objl = -np.log(rp[:, 4])                                         # :344
            if spec.is_focal_loss:
                objl = objl * (np.power(one - rp[:, 4], gamma) * alpha)  # :346
            obj_l = lw[h, 3] * np.sum(objl)                                  # :347
            scale = (dtype(2) - rt[:, 2] * rt[:, 3] / dtype(H * W))[:, None]  # :350
            cint = np.floor(rt[:, 0:2])                                      # :352
            tau = rt[:, 0:2] - cint                                          # :353
            u = rp[:, 0:2] - cint                                            # :354  == (s+c)-c
            xy = -(tau * np.log(u) + (one - tau) * np.log(one - u))          # :355
            xy_l = lw[h, 0] * np.sum(scale * xy)                             # :356
            wh = np.square(np.log(rt[:, 2:4]) - np.log(rp[:, 2:4]))          # :358
            wh_l = lw[h, 1] * np.sum(scale * wh)                             # :359
            if spec.class_num >= 1:
                cls = rt[:, 4].astype(np.int32)                              # :362 cast truncates
                onehot = (cls[:, None] == np.arange(spec.class_num)[None, :]).astype(dtype)
                cls_l = lw[h, 4] * np.sum(-onehot * np.log(rp[:, 5:]))      # :363-364
            else:
                cls_l = dtype(0)                                             # :366
        loss[:, h] = [xy_l, wh_l, noobj_l, obj_l, cls_l]                      # :368
        decisions["heads"].append(dict(
            rows=p["rows"].astype(np.int32), cols=p["cols"].astype(np.int32), best=p["best"],
            rho=p["rho"], own=own[h], object_mask=obj, background_mask=bg, max_iou=p["max_iou"],
            riou=p["riou"],
            margin_best=_margin_sorted(p["riou"]) if V else np.zeros((0,)),
            margin_bg=np.abs(p["max_iou"].astype(np.float64) - np.float64(thresh)),
        ))
    if V:
        r = np.stack(rhos, 0).astype(np.float64)
        m = np.full((V,), np.inf)
        for a in range(3):
            for b in range(a + 1, 3):
                m = np.minimum(m, np.abs(r[a] - r[b]))
        decisions["margin_own"] = m
    else:
        decisions["margin_own"] = np.zeros((0,))
    return loss, decisions


def loss(targets, predicts, spec: HeadSpec, rectified_active=False, dtype=np.float32, want_decisions=True):
    """YOLOv3Loss.loss (:81-138).  Returns (total, detail(6,3) rows [xy,wh,noobj,obj,cls,rect], decisions)."""
    dtype = np.dtype(dtype).type
    dec = decode(predicts, spec, dtype)
    lab = label_decode(targets, spec, dtype)
    n = np.asarray(predicts).shape[0]
    per_image = np.zeros((n, 5, 3), dtype)
    decisions = []
    for i in range(n):                                                       # tf.map_fn, serial (:111)
        l, d = single_image([(dec[h][1][i], dec[h][2][i]) for h in range(3)],
                            [(lab[h][0][i], lab[h][1][i]) for h in range(3)], spec, dtype)
        per_image[i] = l
        if want_decisions:
            decisions.append(d)
    mean = np.sum(per_image, axis=0) / dtype(n)                               # reduce_mean (:112)
    detail = np.zeros((6, 3), dtype)
    detail[:5] = mean
    if rectified_active:                                                     # :125-134, :140-164
        rw = _const(spec.rectified_loss_weight, dtype)
        for h in range(3):
            per_img = np.sum(np.square(dec[h][0]).reshape(n, -1), axis=1)    # :153-155
            detail[5, h] = rw[h] * (np.sum(per_img) / dtype(n))              # :157-162
    total = np.sum(detail[:5]) + (np.sum(detail[5]) if rectified_active else dtype(0))   # :137
    return dtype(total), detail, decisions


# --------------------------------------------------------------------------- analytic backward
def loss_grad(targets, predicts, spec: HeadSpec, rectified_active=False, upstream=1.0, dtype=np.float64):
    """Closed-form d total / d predicts in packed layout: what TF autodiff
    of yolov3_loss.py yields.  No gradient flows through IoU / masks / ownership / targets."""
    dtype = np.dtype(dtype).type
    predicts = np.asarray(predicts)
    n = predicts.shape[0]
    raws = [r.astype(dtype) for r in unpack(predicts, spec)]
    _, _, decisions = loss(targets, predicts, spec, rectified_active, dtype)
    dec = decode(predicts, spec, dtype)
    lab = label_decode(targets, spec, dtype)
    lo, hi = clip_bounds(spec, dtype)
    lw = _const(spec.loss_weights, dtype)
    rw = _const(spec.rectified_loss_weight, dtype)
    kappa = dtype(upstream) / dtype(n)
    gam, alpha = _const(spec.focal_gamma, dtype), _const(spec.focal_alpha, dtype)
    one = dtype(1)
    grads = [np.zeros_like(r) for r in raws]
    for h in range(3):
        H, W = spec.head_grid_sizes[h]
        anchors = spec.anchors_grid(dtype)[h]
        raw, g = raws[h], grads[h]
        s = _sigmoid(raw[..., 0:5])                                          # unclipped sigmoids (t2,t3 unused)
        passes = np.logical_and(s >= lo, s <= hi)
        for i in range(n):
            d = decisions[i]["heads"][h]
            p = dec[h][1][i][..., 4]
            bg = d["background_mask"].astype(dtype)
            if spec.is_focal_loss:
                fprime = np.power(p, gam) / (one - p) - gam * np.power(p, gam - one) * np.log(one - p)
            else:
                fprime = one / (one - p)
            g[i, ..., 4] += kappa * bg * lw[h, 2] * fprime * passes[i, ..., 4] * s[i, ..., 4] * (one - s[i, ..., 4])
            vidx = decisions[i]["valid_index"]
            pos = np.nonzero(d["own"])[0]
            for j in pos:                                                    # duplicates accumulate (T3)
                r, c, b = d["rows"][j], d["cols"][j], d["best"][j]
                tgt = lab[h][0][i][vidx[j]]
                dd = dec[h][1][i][r, c, b]
                pj = dd[4]
                s5 = s[i, r, c, b]
                ps5 = passes[i, r, c, b]
                if spec.is_focal_loss:
                    gprime = alpha * (-np.power(one - pj, gam) / pj + gam * np.power(one - pj, gam - one) * np.log(pj))
                else:
                    gprime = -one / pj
                g[i, r, c, b, 4] += kappa * lw[h, 3] * gprime * ps5[4] * s5[4] * (one - s5[4])
                sc = dtype(2) - tgt[2] * tgt[3] / dtype(H * W)
                k = np.floor(tgt[0:2])
                tau = tgt[0:2] - k
                u = dd[0:2] - k
                g[i, r, c, b, 0:2] += kappa * lw[h, 0] * sc * (-tau / u + (one - tau) / (one - u)) * ps5[0:2] * s5[0:2] * (one - s5[0:2])
                P = dd[2:4]
                g[i, r, c, b, 2:4] += kappa * lw[h, 1] * sc * (-dtype(2) * (np.log(tgt[2:4]) - np.log(P)) / P) * anchors[b] * np.exp(raw[i, r, c, b, 2:4])
                if spec.class_num >= 1:
                    c_id = int(tgt[4])
                    if 0 <= c_id < spec.class_num:
                        z = raw[i, r, c, b, 5:]
                        e = np.exp(z - np.max(z))
                        q = e / np.sum(e)
                        qc = q[c_id]
                        if lo <= qc <= hi:
                            delta = np.zeros_like(q)
                            delta[c_id] = one
                            g[i, r, c, b, 5:] += kappa * lw[h, 4] * (-one / dd[5 + c_id]) * qc * (delta - q)
        if rectified_active:
            g[..., 0:4] += dtype(upstream) * rw[h] * dtype(2) * raw[..., 0:4] / dtype(n)
    return pack_heads(*[g.reshape(n, spec.head_grid_sizes[h][0], spec.head_grid_sizes[h][1], -1)
                        for h, g in enumerate(grads)])
