"""TEST INFRASTRUCTURE ONLY (oracle). Not imported by the product package.

The same graph as oracle/ref_numpy.py, written with torch CPU ops so that
``torch.autograd`` stands in for TF autodiff (the reference has no backward code:
SURVEY.md 8(a) row a18).  ``dtype=torch.float64`` is the ground-truth twin used to
scale tolerances.  Structure follows the reference: one serial pass per image
(tf.map_fn(parallel_iterations=1), yolov3/yolov3_loss.py:111).

Follows yolov3/yolov3_decoder.py:89-192, yolov3/label_decoder.py:26-60,
yolov3/yolov3_loss.py:81-369.  TF semantics mirrored:
``torch.clamp`` passes gradient at both bounds like Minimum/Maximum grads,
``torch.argmax`` = first index, ``amax`` splits ties like reduce_max's gradient,
advanced-index gather backward accumulates duplicates like gather_nd's.

Parity status: **pinned** -- tests/test_oracle_pinned.py holds this file to tests/golden/loss_*.npz, the outputs of the
reference's own yolov3_loss.py / yolov3_decoder.py / label_decoder.py run unmodified under the torch-backed ``tensorflow``
stand-in (oracle/tf_standin, tests/golden/make_loss_golden.py): fp32 forward bit for bit, fp64 twin <= 1e-12.  What stays
assumed about TensorFlow itself is listed in oracle/tf_standin/tensorflow/__init__.py (A1-A6).
Graph constants (eps, 1-eps, anchors, loss weights, iou_thresh, focal alpha/gamma, rectified weights) are rounded to
fp32 once and widened in the fp64 twin: the twin is "the fp32 graph's constants and inputs, exact arithmetic".
Also used, timed, as bench.py's ``cpu_baseline`` / ``--impl reference`` arm.
"""
import torch

from .head_spec import HeadSpec


def unpack(predicts, spec: HeadSpec):
    n = predicts.shape[0]
    out = []
    for h in range(3):
        H, W = spec.head_grid_sizes[h]
        s, e = spec.chan_base[h], spec.chan_base[h] + spec.head_channels[h]
        out.append(predicts[..., s:e].reshape(n, H, W, spec.box_num[h], spec.box_len))
    return out


def _const(x, dt):
    """a graph constant: rounded once to fp32 (ops.convert_to_tensor(v, dtype=float32)), then widened for the twin."""
    return torch.tensor(x, dtype=torch.float32).to(dt)


class _Sigmoid(torch.autograd.Function):
    """tf.sigmoid: 1 / (1 + exp(-x)); registered gradient dy * y * (1 - y) (math_grad._SigmoidGrad)."""

    @staticmethod
    def forward(ctx, x):
        one = torch.ones((), dtype=x.dtype)
        y = one / (one + torch.exp(-x))
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        return dy * (y * (1 - y))


class _Softmax(torch.autograd.Function):
    """tf.nn.softmax: exp(x - max x) * (1 / sum); registered gradient (dy - sum(dy * y)) * y (nn_grad._SoftmaxGrad)."""

    @staticmethod
    def forward(ctx, x):
        e = torch.exp(x - torch.amax(x, dim=-1, keepdim=True))
        y = e * (torch.ones((), dtype=x.dtype) / torch.sum(e, dim=-1, keepdim=True))
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        return (dy - torch.sum(dy * y, dim=-1, keepdim=True)) * y


def _clip(x, lo, hi):
    """tf.clip_by_value = maximum(minimum(x, hi), lo); at equality the gradient goes to x (math_grad._MaximumMinimumGrad)."""
    x = torch.where(x <= hi, x, hi)
    return torch.where(x >= lo, x, lo)


def decode_head(raw, spec: HeadSpec, h):
    H, W = spec.head_grid_sizes[h]
    dt = raw.dtype
    lo = _const(spec.eps, dt)
    hi = _const(1.0 - spec.eps, dt)
    anchors = _const(spec.anchor_boxes[h], dt) * torch.tensor([W, H], dtype=dt)
    ys, xs = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    lt = torch.stack([xs, ys], dim=-1).reshape(H, W, 1, 2).to(dt)
    xy = _clip(_Sigmoid.apply(raw[..., 0:2]), lo, hi) + lt
    wh = torch.exp(raw[..., 2:4]) * anchors
    conf = _clip(_Sigmoid.apply(raw[..., 4:5]), lo, hi)
    parts = [xy, wh, conf]
    if spec.class_num >= 1:
        z = raw[..., 5:]
        z = z - torch.amax(z, dim=-1, keepdim=True)                     # yolov3_decoder.py:189-190
        parts.append(_clip(_Softmax.apply(z), lo, hi))                  # :190-191
    decoded = torch.cat(parts, dim=-1)
    half = wh / 2
    return decoded, torch.cat([xy - half, xy + half], dim=-1)


def decode(predicts, spec: HeadSpec):
    raws = unpack(predicts, spec)
    out = []
    for h in range(3):
        d, b = decode_head(raws[h], spec, h)
        out.append((raws[h][..., 0:4], d, b))
    return out


def label_decode(targets, spec: HeadSpec):
    t = targets.reshape(targets.shape[0], -1, 5)
    out = []
    for h in range(3):
        H, W = spec.head_grid_sizes[h]
        wh_h = torch.tensor([W, H], dtype=t.dtype)
        xy = t[:, :, 0:2] * wh_h
        wh = t[:, :, 2:4] * wh_h
        half = wh / 2
        out.append((torch.cat([xy, wh, t[:, :, 4:5]], dim=-1), torch.cat([xy - half, xy + half], dim=-1)))
    return out


def _iou(pbox, parea, tbox, tarea, tiou):
    lt = torch.maximum(pbox[..., 0:2], tbox[..., 0:2])
    rb = torch.minimum(pbox[..., 2:4], tbox[..., 2:4])
    wh = torch.clamp_min(rb - lt, 0)
    inter = wh[..., 0] * wh[..., 1]
    iou = inter / (parea + tarea - inter)
    if tiou:
        iou = iou * inter / tarea
    return iou


def _single_image(dec, lab, spec: HeadSpec):
    dt = dec[0][0].dtype
    valid = lab[0][0][:, 0] >= 0
    vidx = torch.nonzero(valid)[:, 0]
    V = int(vidx.numel())
    heads = []
    for h in range(3):
        pred, pboxes = dec[h]
        tgt, tb = lab[h][0][vidx], lab[h][1][vidx]
        parea = pred[..., 2] * pred[..., 3]
        grid = torch.floor(tgt[:, 0:2]).to(torch.int64)
        cols, rows = grid[:, 0], grid[:, 1]
        tarea = tgt[:, 2] * tgt[:, 3]
        if V:
            iou = _iou(pboxes[:, :, :, None, :], parea[..., None], tb[None, None, None], tarea, spec.is_tiou_recall)
            max_iou = torch.amax(iou, dim=-1)
            riou = _iou(pboxes[rows, cols], parea[rows, cols], tb[:, None, :], tarea[:, None], spec.is_tiou_recall)
            rho = torch.amax(riou, dim=-1)
            best = torch.argmax(riou, dim=-1)
        else:
            max_iou = torch.full(parea.shape, float("-inf"), dtype=dt)
            rho = torch.zeros((0,), dtype=dt)
            best = torch.zeros((0,), dtype=torch.int64)
        heads.append((tgt, max_iou, rho, rows, cols, best))
    rhos = [hd[2] for hd in heads]
    out = []
    lw = _const(spec.loss_weights, dt)
    gamma, alpha = _const(spec.focal_gamma, dt), _const(spec.focal_alpha, dt)
    one = torch.ones((), dtype=dt)
    for h in range(3):
        tgt, max_iou, rho, rows, cols, best = heads[h]
        H, W = spec.head_grid_sizes[h]
        own = torch.logical_and(rhos[h] >= rhos[(h + 1) % 3], rhos[h] >= rhos[(h + 2) % 3])
        pos = torch.nonzero(own)[:, 0]
        rows, cols, best = rows[pos], cols[pos], best[pos]
        obj = torch.zeros((H, W, spec.box_num[h]), dtype=dt)
        obj[rows, cols, best] = 1.0
        bg = (max_iou < _const(spec.iou_thresh, dt)).to(dt) * (one - obj)
        pred = dec[h][0]
        conf = pred[..., 4]
        noobj = -torch.log(one - conf)
        if spec.is_focal_loss:
            noobj = noobj * torch.pow(conf, gamma)
        noobj_l = lw[h, 2] * torch.sum(noobj * bg)
        rt = tgt[pos]
        rp = pred[rows, cols, best]
        objl = -torch.log(rp[:, 4])
        if spec.is_focal_loss:
            objl = objl * (torch.pow(one - rp[:, 4], gamma) * alpha)
        obj_l = lw[h, 3] * torch.sum(objl)
        scale = (2 - rt[:, 2] * rt[:, 3] / (H * W)).unsqueeze(-1)
        cint = torch.floor(rt[:, 0:2])
        tau = rt[:, 0:2] - cint
        u = rp[:, 0:2] - cint
        xy = -(tau * torch.log(u) + (one - tau) * torch.log(one - u))
        xy_l = lw[h, 0] * torch.sum(scale * xy)
        wh = torch.square(torch.log(rt[:, 2:4]) - torch.log(rp[:, 2:4]))
        wh_l = lw[h, 1] * torch.sum(scale * wh)
        if spec.class_num >= 1:
            cls = rt[:, 4].to(torch.int32)
            onehot = (cls[:, None] == torch.arange(spec.class_num)[None, :]).to(dt)
            cls_l = lw[h, 4] * torch.sum(-onehot * torch.log(rp[:, 5:]))
        else:
            cls_l = torch.zeros((), dtype=dt)
        out.append(torch.stack([xy_l, wh_l, noobj_l, obj_l, cls_l]))
    return torch.stack(out, dim=-1)                                          # (5,3)


def loss(targets, predicts, spec: HeadSpec, rectified_active=False):
    """Returns (total scalar tensor, detail (6,3) tensor)."""
    dec = decode(predicts, spec)
    lab = label_decode(targets, spec)
    n = predicts.shape[0]
    per_image = []
    for i in range(n):
        per_image.append(_single_image([(dec[h][1][i], dec[h][2][i]) for h in range(3)],
                                       [(lab[h][0][i], lab[h][1][i]) for h in range(3)], spec))
    mean = torch.mean(torch.stack(per_image, dim=0), dim=0)
    rect = torch.zeros((1, 3), dtype=predicts.dtype)
    if rectified_active:
        rw = _const(list(spec.rectified_loss_weight), predicts.dtype)
        rect = torch.stack([rw[h] * torch.mean(torch.sum(torch.square(dec[h][0]), dim=[1, 2, 3, 4]), dim=0, keepdim=True)
                            for h in range(3)], dim=-1)
        total = torch.sum(torch.cat([mean, rect], dim=0))
    else:
        total = torch.sum(mean)
    return total, torch.cat([mean, rect], dim=0)


def loss_and_grad(targets, predicts, spec: HeadSpec, rectified_active=False, dtype=torch.float32):
    """forward + autograd backward; returns (total, detail(6,3), grad wrt predicts) as numpy."""
    p = torch.as_tensor(predicts).to(dtype).clone().requires_grad_(True)
    t = torch.as_tensor(targets).to(dtype)
    total, detail = loss(t, p, spec, rectified_active)
    total.backward()
    return total.detach().numpy(), detail.detach().numpy(), p.grad.detach().numpy()
