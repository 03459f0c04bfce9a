"""Generates tests/golden/loss_*.npz and tests/golden/labels_reference.npz by running the REFERENCE's own code

    /root/reference/yolov3/yolov3_loss.py      (YOLOv3Loss.loss and everything it calls, :81-369)
    /root/reference/yolov3/yolov3_decoder.py   (YOLOv3Decoder.decode, :62-192)
    /root/reference/yolov3/label_decoder.py    (LabelDecoder.decode, :26-60)
    /root/reference/dataset/file_util.py       (_parse_string_line :22-32, _parse_image_with_label :35-52)

UNMODIFIED, imported from /root/reference, on top of the torch-backed ``tensorflow`` stand-in of oracle/tf_standin
(TensorFlow itself cannot be installed here).  ``tf.gradients`` (torch autograd) gives the backward.  Each case runs
twice: the fp32 graph, and the fp64 twin (same fp32 constants and inputs, exact arithmetic) that scales tolerances.
Integer decisions (valid rows, centre cells, best anchors, head ownership, object / background masks) are read off the
reference's own intermediate tensors by wrapping the bound methods ``_calc_iou`` / ``_single_head_loss`` of the
instance and tapping the results of its ``sparse_to_dense`` / ``multiply`` calls -- no line of the reference is edited.

Only runnable in the build container (the GPU box has no /root/reference); outputs are committed.

    python tests/golden/make_loss_golden.py            # all cases
    python tests/golden/make_loss_golden.py coco_416   # cases whose name contains the argument
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "oracle", "tf_standin"))
sys.path.insert(0, "/root/reference")

import tensorflow as tf  # noqa: E402  (the stand-in)
from tensorflow import keras  # noqa: E402
from dataset.file_util import FileUtil  # noqa: E402  (reference)
from yolov3.label_decoder import LabelDecoder  # noqa: E402  (reference)
from yolov3.yolov3_decoder import YOLOv3Decoder  # noqa: E402  (reference)
from yolov3.yolov3_loss import YOLOv3Loss  # noqa: E402  (reference)

from helpers import CASES, make_predicts, make_spec, make_targets  # noqa: E402

assert tf.__version__.endswith("standin")
FULL_PREDICTS_LIMIT = 100_000          # floats; larger inputs are regenerated from the seed and checked by hash
GRAD_ENTRY_LIMIT, GRAD_SAMPLE = 60_000, 12_000
DETAIL_VARS = ["coord_loss_xy", "coord_loss_wh", "noobj_iou_loss", "obj_iou_loss", "class_loss", "rectified_coord_loss"]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def build_loss(spec, rectified_coord_num):
    """the call of yolov3/trainer.py:76-83."""
    keras.backend.set_epsilon(spec.eps)                                   # run.py:26
    return YOLOv3Loss(spec.head_grid_sizes, spec.class_num, spec.anchor_boxes, spec.iou_thresh, spec.loss_weights,
                      rectified_coord_num=rectified_coord_num, rectified_loss_weight=list(spec.rectified_loss_weight),
                      is_focal_loss=spec.is_focal_loss, focal_alpha=spec.focal_alpha, focal_gamma=spec.focal_gamma,
                      is_tiou_recall=spec.is_tiou_recall)


def instrument(loss_obj, rec):
    """wrap the instance's bound methods so that the reference's own intermediates are recorded."""
    calc, head = loss_obj._calc_iou, loss_obj._single_head_loss

    def calc_iou(target, target_boxes, predict, predict_boxes, valid_obj_num):
        seen = {}
        with tf.standin.tap(lambda name, value: seen.setdefault(name, value.detach().numpy().copy())):
            out = calc(target, target_boxes, predict, predict_boxes, valid_obj_num)
        riou = seen["arg_max_input"].astype(np.float64)                    # (V, B) IoUs of the centre cell's anchors
        srt = np.sort(riou, axis=-1)
        margin = srt[:, -1] - srt[:, -2] if riou.shape[1] > 1 else np.full(riou.shape[0], np.inf)
        rec["calc"].append(dict(max_iou=out[0].numpy(), rho=out[1].numpy(), xyz=out[2].numpy(),
                                target=target.numpy(), margin_best=margin))
        return out

    def single_head_loss(head_index, predict, target, max_iou, response_max_iou, target_grid_xyz, max_pos, *a):
        seen = {}

        def tap(name, value):
            seen.setdefault(name, value.detach().numpy().copy())           # first call of each kind

        with tf.standin.tap(tap):
            out = head(head_index, predict, target, max_iou, response_max_iou, target_grid_xyz, max_pos, *a)
        rec["head"].append(dict(max_pos=max_pos.numpy()[:, 0], obj=seen["sparse_to_dense"], bg=seen["multiply"]))
        return out

    loss_obj._calc_iou = calc_iou
    loss_obj._single_head_loss = single_head_loss


def run_loss(spec, p, t, rect, precision, per_image_too=True):
    """one session.run of total loss + gradient.  rect: True -> correction term active at this step."""
    tf.standin.set_precision(precision)
    L = build_loss(spec, 10 ** 9 if rect else -1)
    rec = dict(calc=[], head=[])
    instrument(L, rec)
    per_image = []
    single = L._single_image_loss

    def single_image_loss(item):
        out = single(item)
        per_image.append(out.numpy())
        return out

    L._single_image_loss = single_image_loss
    P, T = tf.standin.input(p, requires_grad=True), tf.standin.input(t)
    total = L.loss(T, P)
    (grad,) = tf.gradients(total, [P])
    tf.standin.end_step()
    detail = np.stack([getattr(L, v).numpy() for v in
                       ["coord_loss_xy", "coord_loss_wh", "noobj_iou_loss", "obj_iou_loss", "class_loss",
                        "rectified_coord_loss"]])
    return dict(total=total.numpy(), detail=detail, per_image=np.stack(per_image), grad=grad.numpy(), rec=rec,
                current_num=int(L.current_num.numpy()))


def decisions_arrays(spec, rec, n, t):
    """flatten what the reference decided, image by image (ragged over valid labels)."""
    m = t.shape[1] // 5
    offs = [0]
    out = dict(valid_index=[], rows=[[], [], []], cols=[[], [], []], best=[[], [], []], rho=[[], [], []],
               own=[[], [], []], margin_best=[[], [], []])
    masks = {f"{k}{h}": [] for k in ("obj", "bg", "maxiou") for h in range(3)}
    H8, W8 = spec.head_grid_sizes[0]
    for i in range(n):
        # valid rows: the reference decides on head /8's x (yolov3_loss.py:239); read back from its gathered targets
        tgt8 = rec["calc"][3 * i]["target"]
        x8 = (t[i].reshape(m, 5)[:, 0].astype(np.float32) * np.float32(W8))
        vidx = np.nonzero(x8 >= 0)[0]
        assert len(vidx) == tgt8.shape[0] and (tgt8.dtype != np.float32 or np.array_equal(x8[vidx], tgt8[:, 0]))
        out["valid_index"].append(vidx)
        offs.append(offs[-1] + len(vidx))
        for h in range(3):
            c, hd = rec["calc"][3 * i + h], rec["head"][3 * i + h]
            xyz = c["xyz"].reshape(-1, 3)
            own = np.zeros(len(vidx), bool)
            own[hd["max_pos"]] = True
            out["rows"][h].append(xyz[:, 0])
            out["cols"][h].append(xyz[:, 1])
            out["best"][h].append(xyz[:, 2])
            out["rho"][h].append(c["rho"])
            out["own"][h].append(own)
            out["margin_best"][h].append(c["margin_best"])
            masks[f"obj{h}"].append(hd["obj"] != 0)
            masks[f"bg{h}"].append(hd["bg"] != 0)
            masks[f"maxiou{h}"].append(c["max_iou"])
    flat = dict(dec_offsets=np.asarray(offs, np.int32),
                dec_valid_index=np.concatenate(out["valid_index"]).astype(np.int32))
    for k in ("rows", "cols", "best"):
        flat["dec_" + k] = np.stack([np.concatenate(out[k][h]) for h in range(3)]).astype(np.int32)
    flat["dec_rho"] = np.stack([np.concatenate(out["rho"][h]) for h in range(3)])
    flat["dec_own"] = np.stack([np.concatenate(out["own"][h]) for h in range(3)])
    flat["dec_margin_best"] = np.stack([np.concatenate(out["margin_best"][h]) for h in range(3)])
    for h in range(3):
        flat[f"obj_bits{h}"] = np.packbits(np.stack(masks[f"obj{h}"]).reshape(-1))
        flat[f"bg_bits{h}"] = np.packbits(np.stack(masks[f"bg{h}"]).reshape(-1))
        flat[f"maxiou{h}"] = np.stack(masks[f"maxiou{h}"])
    return flat


def owner_records(spec, rec, n):
    """packed-layout record numbers (flat over the batch) of the anchors the reference's object masks mark."""
    from oracle.ref_numpy import packed_offset
    out = []
    for i in range(n):
        for h in range(3):
            for (y, x, b) in np.argwhere(rec["head"][3 * i + h]["obj"] != 0):
                out.append(packed_offset(spec, i, h, int(y), int(x), int(b), 0) // spec.box_len)
    return np.asarray(sorted(set(out)), np.int64)


def run_decode(spec, p, t, precision, sample):
    """YOLOv3Decoder.decode + LabelDecoder.decode as run.py:62 / yolov3_loss.py:99,103 call them."""
    tf.standin.set_precision(precision)
    keras.backend.set_epsilon(spec.eps)
    dec = YOLOv3Decoder(spec.head_grid_sizes, spec.class_num, spec.anchor_boxes).decode(tf.standin.input(p))
    lab = LabelDecoder(spec.head_grid_sizes).decode(tf.standin.input(t))
    out = {}
    for h in range(3):
        raw, d, b = (x.numpy() for x in dec[h])
        L = d.shape[-1]
        idx = sample[h]
        out[f"raw{h}"] = raw.reshape(-1, 4)[idx]
        out[f"decoded{h}"] = d.reshape(-1, L)[idx]
        out[f"boxes{h}"] = b.reshape(-1, 4)[idx]
        out[f"lab{h}"] = lab[h][0].numpy()
        out[f"labbox{h}"] = lab[h][1].numpy()
    return out


def save_case(name, spec_name, p, t, rects, spec=None, seeds=None, note="", sample_floats=20000, spec_overrides=None):
    spec = spec or make_spec(spec_name)
    n = p.shape[0]
    H, W, C, anchors, kw = CASES[spec_name]
    meta = dict(case=name, spec_name=spec_name, n=n, m=t.shape[1] // 5, rects=[int(r) for r in rects], note=note,
                seeds=seeds, spec_overrides=spec_overrides or {}, eps=spec.eps,
                generator="tests/golden/make_loss_golden.py", assumptions="oracle/tf_standin/tensorflow/__init__.py A1-A6")
    save = dict(targets=t, predicts_sha256=np.array(sha(p)))
    if p.size <= FULL_PREDICTS_LIMIT:
        save["predicts"] = p
    rng = np.random.default_rng(99)
    sample = []
    for h in range(3):
        cnt = n * spec.head_grid_sizes[h][0] * spec.head_grid_sizes[h][1] * spec.box_num[h]
        sample_records = max(sample_floats // spec.box_len, 64)
        sample.append(np.arange(cnt) if cnt <= sample_records else np.sort(rng.choice(cnt, sample_records, replace=False)))
        save[f"sample{h}"] = sample[h].astype(np.int32)
    for prec, tag in (("float32", "32"), ("float64", "64")):
        d = run_decode(spec, p, t, prec, sample)
        for k, v in d.items():
            if tag == "64" and (k.startswith("raw") or k.startswith("lab")):
                continue                                   # raw logits and label decode: fp32 is the contract
            save[f"{k}_{tag}"] = v
    for rect in rects:
        r32 = run_loss(spec, p, t, rect, "float32")
        r64 = run_loss(spec, p, t, rect, "float64")
        sfx = f"_rect{int(rect)}"
        g32, g64 = r32["grad"].reshape(-1), r64["grad"].reshape(-1)
        with np.errstate(invalid="ignore"):
            nz = (g32 != 0) | (g64 != 0)                          # NaN != 0 is True: non-finite entries are kept
        complete = True
        if rect and nz.sum() > GRAD_ENTRY_LIMIT:
            # the correction term touches t0..t3 of EVERY anchor (2 rw t / N): keep every entry of the records that own
            # a label, every conf / class entry, and a random sample of the rest
            L = spec.box_len
            rec_nz = nz.reshape(-1, L)
            keep = np.zeros_like(rec_nz)
            keep[:, 4:] = rec_nz[:, 4:]
            owners = owner_records(spec, r32["rec"], n)
            keep[owners] = rec_nz[owners]
            rest = np.nonzero((rec_nz & ~keep).reshape(-1))[0]
            keep = keep.reshape(-1)
            keep[np.random.default_rng(7).choice(rest, min(GRAD_SAMPLE, len(rest)), replace=False)] = True
            nz, complete = keep, False
        idx = np.nonzero(nz)[0]
        meta[f"grad_complete{sfx}"] = bool(complete)
        save["grad_idx" + sfx] = idx.astype(np.int64 if p.size >= 2 ** 31 else np.int32)
        save["grad32" + sfx] = g32[idx]
        save["grad64" + sfx] = g64[idx]
        for tag, r in (("32", r32), ("64", r64)):
            save[f"total{tag}{sfx}"] = r["total"]
            save[f"detail{tag}{sfx}"] = r["detail"]
            save[f"per_image{tag}{sfx}"] = r["per_image"]
        if rect == rects[0]:
            save.update(decisions_arrays(spec, r32["rec"], n, t))
            d32 = decisions_arrays(spec, r32["rec"], n, t)
            d64 = decisions_arrays(spec, r64["rec"], n, t)
            save["dec_rho64"], save["dec_margin_best"] = d64["dec_rho"], d64["dec_margin_best"]     # margins from the fp64 twin
            # does the fp64 twin take the same discrete decisions?  (a label centre within an fp32 ulp of a cell border
            # lands in another cell in fp64: then only the fp32 run is a valid target for VALUES)
            meta["twin_agrees"] = bool(all(np.array_equal(d32[k], d64[k]) for k in
                                           ("dec_rows", "dec_cols", "dec_best", "dec_own", "obj_bits0", "obj_bits1", "obj_bits2",
                                            "bg_bits0", "bg_bits1", "bg_bits2")))
        meta[f"current_num_after{sfx}"] = r32["current_num"]
    save["meta"] = np.array(json.dumps(meta))
    path = os.path.join(HERE, f"loss_{name}.npz")
    np.savez_compressed(path, **save)
    print("%-34s n=%d total32=%s  nnz grad=%d  %.0f KB  twin_agrees=%s" % (
        name, n, [float(save[f"total32_rect{int(r)}"]) for r in rects], len(idx), os.path.getsize(path) / 1024,
        meta["twin_agrees"]))
    return meta


# ------------------------------------------------------------------------------------------------ label pipeline
def reference_labels(image_size):
    """label.txt lines of dataset/test_sample through the reference's parser and letter-box transform."""
    src = "/root/reference/dataset/test_sample"
    tf.standin.set_precision("float32")
    out, shapes, lines = [], [], []
    for line in open(os.path.join(src, "label.txt")):
        if not line.strip():
            continue
        path, labels = FileUtil._parse_string_line(line.rstrip("\n"), os.path.join(src, "images"))
        image, lab = FileUtil._parse_image_with_label(path, labels, tf.constant(image_size, dtype=tf.int32))
        assert image.shape[:2] == tuple(image_size)
        import cv2
        shapes.append(cv2.imread(path).shape[:2])
        out.append(lab.numpy())
        lines.append(line.rstrip("\n"))
    return out, shapes, lines


def padded_batch(labels):
    """tf.data padded_batch(batch, TensorShape([None]), padding_values=-1.) of dataset/file_util.py:97: every row padded
    with -1 to the longest row of the batch (tf.data itself is not re-hosted; this is its documented behaviour)."""
    m = max(len(l) for l in labels)
    out = np.full((len(labels), m), -1.0, np.float32)
    for i, l in enumerate(labels):
        out[i, :len(l)] = l
    return out


def save_labels():
    save = {}
    for tag, size in (("416", (416, 416)), ("384x480", (384, 480))):
        labs, shapes, lines = reference_labels(size)
        save["flat_" + tag] = np.concatenate(labs)
        save["batch_" + tag] = padded_batch(labs)
    save["offsets"] = np.cumsum([0] + [len(l) // 5 for l in labs]).astype(np.int32)
    save["shapes"] = np.asarray(shapes, np.int32)
    save["lines"] = np.array(lines)
    np.savez_compressed(os.path.join(HERE, "labels_reference.npz"), **save)
    print("labels_reference: %d images, %d boxes" % (len(labs), save["offsets"][-1]))
    return save


# ------------------------------------------------------------------------------------------------ cases
def std_inputs(spec_name, n, m, seed):
    """the inputs tests/test_gpu_parity.py::_inputs builds."""
    spec = make_spec(spec_name)
    return make_predicts(spec, n, seed=seed), make_targets(spec, n, m, seed=seed + 1, min_valid=0, pad_in_middle=True)


def edge_labels(t, n, m, spec):
    """V = 0 image, a duplicated label (rule T3), a class id >= C: the edge rows of SURVEY 8(c)."""
    t = t.reshape(n, m, 5).copy()
    if n >= 3:
        t[1] = -1.0
    last = n - 1
    v = np.nonzero(t[last, :, 0] >= 0)[0]
    if len(v) >= 2:
        t[last, v[1]] = t[last, v[0]]
    if len(v) >= 3:
        t[last, v[2], 4] = spec.class_num + 2
    return t.reshape(n, m * 5)


def border_labels(spec, n, m, seed):
    """labels whose centres sit exactly on cell borders of the three grids (fl32(k / W) * W is k or just below k)."""
    rng = np.random.default_rng(seed)
    t = np.full((n, m, 5), -1.0, np.float32)
    (H8, W8), (H16, W16), (H32, W32) = spec.head_grid_sizes

    def same_cell_in_both(k, W):
        """fl32(k / W) * W lands in the same cell of all three grids whether the product is rounded to fp32 or not
        (the other kind -- fp32 says cell k, exact arithmetic k - 1 -- is the demo set's x = 232/480 label)"""
        v = np.float32(k / W)
        return all(np.floor(np.float32(v * np.float32(g))) == np.floor(np.float64(v) * g) for g in (W8, W16, W32))

    for i in range(n):
        for j in range(m - 2):
            W, H = [(W8, H8), (W16, H16), (W32, H32)][j % 3]
            while True:
                kx, ky = int(rng.integers(1, W)), int(rng.integers(1, H))
                if same_cell_in_both(kx, W) and same_cell_in_both(ky, H):
                    break
            w, h = rng.uniform(0.03, 0.3, 2)
            t[i, j] = [np.float32(kx / W), np.float32(ky / H), w, h, int(rng.integers(0, max(spec.class_num, 1)))]
        k = W8 - 1
        while not same_cell_in_both(k, W8):
            k -= 1
        t[i, m - 2] = [np.float32(k / W8), np.float32(k / H8), 0.02, 0.02, 0]     # the last /8 columns (c = 75 at 608)
    return t.reshape(n, m * 5)


def main(only=None):
    def want(name):
        return only is None or only in name

    if want("labels"):
        labels = save_labels()
    else:
        labels = np.load(os.path.join(HERE, "labels_reference.npz"))

    # the small shapes of tests/test_gpu_parity.py, rectified off and on, edge rows in
    for spec_name in ["tiny_32_c2", "coco_64x96_c4", "coco_128_c3_focal_tiou", "single_160_c1_focal", "coco_96x64_c13",
                      "coco_64_c20"]:
        if want(spec_name):
            spec = make_spec(spec_name)
            p, t = std_inputs(spec_name, 3, 8, 11)
            save_case(spec_name, spec_name, p, edge_labels(t, 3, 8, spec), [False, True], seeds=dict(predicts=11, targets=12),
                      note="std inputs (seed 11), image 1 emptied, a duplicate label and a class id >= C in image 2")

    # BASELINE configs[0]: the demo labels, batch 4, as shipped (384x480, 3-2-3 anchors, class_num 0, rectified window open)
    if want("demo_384x480"):
        spec = make_spec("shipped_384x480_c0")
        t = padded_batch([labels["flat_384x480"][labels["offsets"][i] * 5:labels["offsets"][i + 1] * 5] for i in (3, 4, 15, 16)])
        p = make_predicts(spec, 4, seed=21)
        save_case("demo_384x480_c0", "shipped_384x480_c0", p, t, [True, False], seeds=dict(predicts=21),
                  note="dataset/test_sample images 4,5,16,17 through file_util.py:47-52 at 384x480; image 17 holds the "
                       "x = 232/480 label whose fp32 product with 60 is exactly 29")
    # ... and as named (416x416); the demo classes are 0..12
    if want("demo_416"):
        spec = make_spec("demo_416_c13")
        t = padded_batch([labels["flat_416"][labels["offsets"][i] * 5:labels["offsets"][i + 1] * 5] for i in (0, 3, 4, 16)])
        p = make_predicts(spec, 4, seed=22)
        save_case("demo_416_c13", "demo_416_c13", p, t, [False], seeds=dict(predicts=22),
                  note="dataset/test_sample images 1,4,5,17 letter-boxed to 416x416, 13 classes")

    # BASELINE configs[1]: 416^2 / 80 classes
    if want("coco_416_c80"):
        spec = make_spec("coco_416_c80")
        p, t = std_inputs("coco_416_c80", 2, 16, 31)
        save_case("coco_416_c80", "coco_416_c80", p, edge_labels(t, 2, 16, spec), [False], seeds=dict(predicts=31, targets=32),
                  note="std inputs (seed 31), duplicate label and class id >= C in image 1")
    # BASELINE configs[2]: 608^2 / 80 classes; labels on cell borders up to column 75 (SURVEY F8's worst case)
    if want("coco_608_c80"):
        spec = make_spec("coco_608_c80")
        p = make_predicts(spec, 2, seed=41)
        meta = save_case("coco_608_c80", "coco_608_c80", p, border_labels(spec, 2, 12, 42), [False],
                         seeds=dict(predicts=41, targets=42),
                         note="centres exactly on /8, /16, /32 cell borders; one label in the last /8 columns (c = 75)")
        assert meta["twin_agrees"]
    # BASELINE configs[4]: 512^2, class_num 0, three anchors per head, focal + rectified
    if want("single_512"):
        spec = make_spec("single_512_c0_focal")
        p, t = std_inputs("single_512_c0_focal", 2, 16, 51)
        save_case("single_512_c0_focal", "single_512_c0_focal", p, t, [True], seeds=dict(predicts=51, targets=52),
                  note="focal loss (alpha 1, gamma 2) and the prior-box correction term active")


    # IoU exactly at iou_thresh: zero box logits + power-of-two anchors make every IoU an exactly rounded
    # quotient; label (w, h) = (3, 4) cells on a cell centre has IoU 8/12 with the (2, 4) anchor of its own cell
    if want("pow2_128_c2"):
        spec = make_spec("pow2_128_c2")
        p = make_predicts(spec, 2, seed=71)
        heads = [h.copy() for h in __import__("oracle.ref_numpy", fromlist=["unpack"]).unpack(p, spec)]
        from oracle.ref_numpy import pack_heads
        for h in heads:
            h[..., 0:4] = 0.0
        p = pack_heads(*[h.reshape(h.shape[0], h.shape[1], h.shape[2], -1) for h in heads])
        t = np.full((2, 6, 5), -1.0, np.float32)
        t[0, 0] = [5.5 / 16, 7.5 / 16, 3.0 / 16, 4.0 / 16, 1]       # /8 head: centre of cell (7, 5), 3 x 4 cells
        t[0, 2] = [10.5 / 16, 3.5 / 16, 2.0 / 16, 2.0 / 16, 0]      # IoU 0.25 / 0.5 / 0.5 in its cell: anchors 1, 2 tie (T1)
        t[1, 1] = [2.5 / 8, 2.5 / 8, 3.0 / 8, 4.0 / 8, 1]           # the same on the /16 grid
        t[1, 4] = [0.5, 0.5, 1.0 / 16, 1.0 / 16, 0]                 # centre on a border of all three grids
        t = t.reshape(2, 30)
        save_case("pow2_128_c2", "pow2_128_c2", p, t, [False], seeds=None, sample_floats=4000,
                  note="t0..t3 = 0 everywhere; iou_thresh = fl32(8/12) is met exactly by non-owner anchors")
        z = np.load(os.path.join(HERE, "loss_pow2_128_c2.npz"))
        thr = np.float32(spec.iou_thresh)
        hits = sum(int(np.sum((z[f"maxiou{h}"] == thr))) for h in range(3))
        assert hits >= 2, "no anchor sits exactly on the threshold"
        print("   anchors with max IoU == iou_thresh exactly:", hits)

    # saturated logits: where the reference produces inf / NaN
    if want("saturated"):
        spec = make_spec("coco_64x96_c4")
        p, t = std_inputs("coco_64x96_c4", 2, 4, 61)
        t = t.reshape(2, 4, 5)
        t[1, 0] = [0.30, 0.40, 0.2, 0.3, 1]
        t[1, 1] = [0.70, 0.60, 0.1, 0.1, 2]
        t[1, 2:] = -1
        from oracle.ref_numpy import pack_heads, unpack
        heads = [h.copy() for h in unpack(p, spec)]
        for h in range(3):
            H, W = spec.head_grid_sizes[h]
            cx, cy = int(np.float32(0.30) * np.float32(W)), int(np.float32(0.40) * np.float32(H))
            heads[h][1, cy, cx, :, 0] = 20.0            # sigmoid(t0) = 1.0 in fp32: (s + c) - c = 1, log(1 - u) = -inf
            cx, cy = int(np.float32(0.70) * np.float32(W)), int(np.float32(0.60) * np.float32(H))
            heads[h][1, cy, cx, :, 5 + 2] = -200.0      # softmax prob of the label's class underflows: clipped to eps
        heads[0][1, 0, 0, 0, 4] = 20.0                  # a background anchor with conf = 1.0: -log(1 - 1) = inf
        heads[0][1, 0, 1, 0, 4] = -20.0                 # conf below eps: clipped, gradient gated off
        heads[1][1, 0, 0, 1, 2] = 100.0                 # exp overflows: an infinite box far from every label
        p = pack_heads(*[h.reshape(h.shape[0], h.shape[1], h.shape[2], -1) for h in heads])
        with np.errstate(all="ignore"):
            save_case("saturated_64x96_c4", "coco_64x96_c4", p, t.reshape(2, 20), [False], seeds=None,
                      note="image 0 finite; image 1 holds saturated x, conf, class and w logits")

    # the tf.cond counter of yolov3_loss.py:125-134,152 over consecutive steps
    if want("rect_counter"):
        spec = make_spec("tiny_32_c2")
        p, t = std_inputs("tiny_32_c2", 3, 4, 11)
        tf.standin.set_precision("float32")
        L = build_loss(spec, 3)
        cur, rect_rows, totals = [], [], []
        for step in range(4):
            total = L.loss(tf.standin.input(t), tf.standin.input(p))
            tf.standin.end_step()
            cur.append(int(L.current_num.numpy()))
            rect_rows.append(L.rectified_coord_loss.numpy())
            totals.append(total.numpy())
        np.savez_compressed(os.path.join(HERE, "seq_rect_counter.npz"), predicts=p, targets=t, rectified_coord_num=3,
                            current_num=np.asarray(cur), rect_rows=np.stack(rect_rows), totals=np.asarray(totals))
        print("rect_counter: current_num", cur, "active", [bool(np.abs(r).sum() > 0) for r in rect_rows])


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else None)
