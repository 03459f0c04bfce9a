"""Drop-in for yolov3/yolov3_post_process.py::YOLOv3PostProcessor backed by yv3_filter_* / yv3_nms.

Two surfaces:

* ``YOLOv3PostProcessor`` -- the reference's static methods with the reference's NumPy in/out contract
  (run.py:69-77 calls them per image).  Arrays go to the GPU, the kernels do the filtering / NMS, the
  reference's row format (and its id-collision quirk) is rebuilt around the result.
* ``HeadPostProcessor`` -- the batched, device-resident path: packed logits of N images in, candidate
  rows + per-head counts + survivor flags out, nothing materialised in between.
"""
import ctypes

import numpy as np
import torch

from . import _lib

_NEUTRAL = (1.0, 1.0)


def _f32_exact(a):
    a = np.asarray(a)
    if a.dtype == np.float32:
        return a
    b = a.astype(np.float32)
    if not np.array_equal(b.astype(a.dtype), a):
        raise ValueError("candidate rows must be exactly representable in fp32 (as filter_boxes produces "
                         "them): the GPU path stores rows as fp32 and evaluates IoU in fp64 from those")
    return b


class _NmsRunner(object):
    """yv3_nms on one or more candidate lists already on the device."""

    @staticmethod
    def run(cand, counts, num_classes, nms_thresh, cfg=None):
        n, capacity = cand.shape[0], cand.shape[1]
        lib = _lib.lib()
        size = ctypes.c_size_t(0)
        cfg_ref = ctypes.byref(cfg) if cfg is not None else None
        _lib.check(lib.yv3_query_postprocess_workspace(cfg_ref, n, capacity, ctypes.byref(size)),
                   "yv3_query_postprocess_workspace")
        ws = torch.empty(size.value, dtype=torch.uint8, device=cand.device)
        keep = torch.empty((n, capacity), dtype=torch.uint8, device=cand.device)
        rc = lib.yv3_nms(cfg_ref, _lib.ptr(cand), _lib.ptr(counts), counts.shape[1], n, capacity,
                         int(num_classes), float(nms_thresh), _lib.ptr(keep), _lib.ptr(ws), ws.numel(),
                         _lib.current_stream_ptr(cand.device))
        _lib.check(rc, "yv3_nms")
        return keep


class YOLOv3PostProcessor(object):
    HEAD_BOX_COLOR = [[255, 0, 0], [0, 255, 0], [0, 0, 255]]
    _cfg_cache = {}

    @staticmethod
    def _cfg_for(shapes):
        key = tuple(shapes)
        cfg = YOLOv3PostProcessor._cfg_cache.get(key)
        if cfg is None:
            grids = [(s[0], s[1]) for s in shapes]
            anchors = [[_NEUTRAL] * s[2] for s in shapes]
            cfg = _lib.make_config(grids, shapes[0][3] - 5, anchors)
            YOLOv3PostProcessor._cfg_cache[key] = cfg
        return cfg

    @staticmethod
    def filter_boxes(head_8_prediction, head_8_boxes, head_16_prediction, head_16_boxes,
                     head_32_prediction, head_32_boxes, score_thresh):
        """yolov3_post_process.py:21-76.  Inputs: decoded heads (H,W,B,5+C) and corner boxes (H,W,B,4) of
        ONE image (NumPy or tensors).  Returns 3 arrays (k_h, 8): float64 when C >= 1 (the reference's
        rows are promoted by the int64 arg-max column), the input dtype when C == 0, and
        ``np.empty((0, 8), float64)`` for a head without survivors (:64)."""
        preds = [head_8_prediction, head_16_prediction, head_32_prediction]
        boxes = [head_8_boxes, head_16_boxes, head_32_boxes]
        shapes = [tuple(int(v) for v in p.shape) for p in preds]
        if any(len(s) != 4 for s in shapes) or len({s[3] for s in shapes}) != 1:
            raise ValueError("each head prediction must be (H, W, B, 5 + class_num)")
        cfg = YOLOv3PostProcessor._cfg_for(shapes)
        dp = [_lib.to_device_f32(p) for p in preds]
        db = [_lib.to_device_f32(b) for b in boxes]
        device = dp[0].device
        lib = _lib.lib()
        capacity = int(sum(s[0] * s[1] * s[2] for s in shapes))
        size = ctypes.c_size_t(0)
        _lib.check(lib.yv3_query_postprocess_workspace(ctypes.byref(cfg), 1, capacity, ctypes.byref(size)),
                   "yv3_query_postprocess_workspace")
        ws = torch.empty(size.value, dtype=torch.uint8, device=device)
        cand = torch.empty((capacity, 8), dtype=torch.float32, device=device)
        counts = torch.empty(3, dtype=torch.int32, device=device)
        pp = (ctypes.c_void_p * 3)(*[_lib.ptr(t) for t in dp])
        bp = (ctypes.c_void_p * 3)(*[_lib.ptr(t) for t in db])
        # np.where(all_score > score_thresh): the Python float is compared in the array's dtype (fp32)
        thr = float(np.float32(score_thresh))
        rc = lib.yv3_filter_decoded(ctypes.byref(cfg), pp, bp, thr, _lib.ptr(cand), capacity, _lib.ptr(counts),
                                    _lib.ptr(ws), ws.numel(), _lib.current_stream_ptr(device))
        _lib.check(rc, "yv3_filter_decoded")
        k = counts.cpu().numpy().astype(np.int64)
        rows = cand[: int(k.sum())].cpu().numpy()
        out_dtype = np.float64 if shapes[0][3] > 5 else np.float32
        out, start = [], 0
        for h in range(3):
            if k[h] == 0:
                out.append(np.empty(shape=(0, 8), dtype=np.float64))
            else:
                out.append(rows[start:start + k[h]].astype(out_dtype))
            start += k[h]
        return out

    @staticmethod
    def true_keep_mask(boxes, nms_thresh):
        """Greedy-NMS survivors (a bool mask over the concatenation of the three heads' rows):
        what _apply_nms (:108-129) keeps, before apply_nms' id bookkeeping."""
        lens = [len(b) for b in boxes]
        total = int(sum(lens))
        if total == 0:
            return np.zeros((0,), dtype=bool)
        if total > _lib.YV3_MAX_CANDIDATES:
            raise ValueError("more than %d candidates in one list" % _lib.YV3_MAX_CANDIDATES)
        allb = np.concatenate([np.asarray(b, dtype=np.float64).reshape(-1, np.asarray(b).shape[-1])[:, :8]
                               for b in boxes if len(b)], axis=0)
        rows = _f32_exact(allb)
        cls = rows[:, 6]
        if np.any(cls < 0) or np.any(cls != np.floor(cls)) or np.any(cls >= 16384):
            raise ValueError("class ids must be integers in [0, 16384)")
        cand = _lib.to_device_f32(rows).view(1, total, 8)
        counts = torch.tensor([lens], dtype=torch.int32).to(cand.device)
        keep = _NmsRunner.run(cand, counts, int(cls.max()) + 1, nms_thresh)
        return keep[0].cpu().numpy().astype(bool)

    @staticmethod
    def apply_nms(boxes, nms_thresh, bug_compat=True):
        """yolov3_post_process.py:79-105.  ``boxes``: the 3 arrays of filter_boxes; modified in place and
        returned like the reference does: an empty head becomes a (0, 9) array, any other head a list of
        (9,) rows (the 8 columns + the id column).

        ``bug_compat=True`` reproduces the shipped behaviour bit for bit: ``start_index`` is never
        advanced (:81-89), so ids restart at 0 in every head and a row is returned iff its LOCAL id equals
        the local id of ANY greedy survivor of ANY head (:94-104).  ``bug_compat=False`` returns exactly
        the greedy survivors."""
        lens = [len(b) for b in boxes]
        keep = YOLOv3PostProcessor.true_keep_mask(boxes, nms_thresh)
        offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
        kept_local = set()
        for h in range(3):
            kept_local.update(np.nonzero(keep[offsets[h]:offsets[h + 1]])[0].tolist())
        for h, head_boxes in enumerate(boxes):
            if lens[h] == 0:
                boxes[h] = np.reshape(head_boxes, (0, 9))
                continue
            ids = np.expand_dims(np.arange(0, lens[h], dtype=float), axis=-1)
            rows = np.concatenate([np.asarray(head_boxes), ids], axis=-1)
            if bug_compat:
                sel = [i for i in range(lens[h]) if i in kept_local]
            else:
                sel = np.nonzero(keep[offsets[h]:offsets[h + 1]])[0].tolist()
            boxes[h] = [rows[i] for i in sel]
        return boxes

    @staticmethod
    def resize_boxes(boxes, target_size):
        """yolov3_post_process.py:162-172: box[:4] * target_size on the (few) surviving host rows."""
        return [head_boxes if len(head_boxes) == 0 else
                [np.concatenate([box[:4] * target_size, box[4:]], axis=-1) for box in head_boxes]
                for head_boxes in boxes]


class HeadPostProcessor(object):
    """Batched decode + score threshold + class-aware NMS from packed logits, device resident.

    ``run(predicts, score_thresh, nms_thresh)`` -> dict(cand (N,cap,8) fp32, counts (N,3) int32,
    keep (N,cap) uint8), all CUDA tensors; per image the rows are in the reference's order and ``keep``
    marks the greedy survivors.  ``to_reference(result, i)`` rebuilds run.py's per-image lists."""

    def __init__(self, head_grid_sizes, class_num, anchor_boxes, capacity=None):
        self.head_grid_sizes = [(int(h), int(w)) for (h, w) in head_grid_sizes]
        self.class_num = int(class_num)
        self._cfg = _lib.make_config(self.head_grid_sizes, self.class_num, anchor_boxes)
        self._lib = _lib.lib()
        self.records_per_image = int(self._lib.yv3_records_per_image(ctypes.byref(self._cfg)))
        self.packed_floats_per_image = int(self._lib.yv3_packed_floats_per_image(ctypes.byref(self._cfg)))
        self.capacity = int(min(self.records_per_image if capacity is None else capacity,
                                _lib.YV3_MAX_CANDIDATES))
        self._ws = {}
        self._last_ws = None

    def _workspace(self, n, device):
        key = (n, device)
        ws = self._ws.get(key)
        if ws is None:
            size = ctypes.c_size_t(0)
            _lib.check(self._lib.yv3_query_postprocess_workspace(ctypes.byref(self._cfg), n, self.capacity,
                                                                 ctypes.byref(size)),
                       "yv3_query_postprocess_workspace")
            ws = torch.empty(size.value, dtype=torch.uint8, device=device)
            self._ws[key] = ws
        return ws

    def filter(self, predicts, score_thresh, out=None, sweep_done=None):
        """cand (N, capacity, 8), counts (N, 3) from the packed logits.  ``sweep_done``: an event recorded between the
        HBM-bound score sweep and the small gather kernel (a scheduler's hook, see bench.py --serialize-sweeps)."""
        p = _lib.to_device_f32(predicts)
        if p.dim() != 4 or p[0].numel() != self.packed_floats_per_image:
            raise ValueError("predicts does not have the packed head layout of this configuration")
        n = p.shape[0]
        ws = self._workspace(n, p.device)
        if out is None:
            cand = torch.empty((n, self.capacity, 8), dtype=torch.float32, device=p.device)
            counts = torch.empty((n, 3), dtype=torch.int32, device=p.device)
        else:
            cand, counts = out
        for phases in ((3,) if sweep_done is None else (1, 2)):
            rc = self._lib.yv3_filter_packed(ctypes.byref(self._cfg), _lib.ptr(p), n,
                                             float(np.float32(score_thresh)), _lib.ptr(cand), self.capacity,
                                             _lib.ptr(counts), _lib.ptr(ws), ws.numel(), phases,
                                             _lib.current_stream_ptr(p.device))
            _lib.check(rc, "yv3_filter_packed")
            if sweep_done is not None and phases == 1:
                sweep_done.record(torch.cuda.current_stream(p.device))
        self._last_ws = ws
        return cand, counts

    def filter_heads(self, head_predicts, score_thresh, out=None):
        """``filter`` on the three head tensors before yolov3_detector.py:80-85 packs them ([/8, /16, /32], each
        (N, H_h, W_h, B_h * (5 + C)) fp32 contiguous): same rows, order and counts, without the concatenate copy."""
        if len(head_predicts) != 3:
            raise ValueError("head_predicts must hold the /8, /16 and /32 head tensors")
        heads = [_lib.to_device_f32(p) for p in head_predicts]
        n = heads[0].shape[0]
        for h, p in enumerate(heads):
            gh, gw = self.head_grid_sizes[h]
            if p.shape[0] != n or tuple(p.shape[1:3]) != (gh, gw) or p[0].numel() * 1 != gh * gw * int(self._cfg.box_num[h]) * (5 + self.class_num):
                raise ValueError("head %d has shape %s, expected (N, %d, %d, %d)" % (
                    h, tuple(p.shape), gh, gw, int(self._cfg.box_num[h]) * (5 + self.class_num)))
        ws = self._workspace(n, heads[0].device)
        if out is None:
            cand = torch.empty((n, self.capacity, 8), dtype=torch.float32, device=heads[0].device)
            counts = torch.empty((n, 3), dtype=torch.int32, device=heads[0].device)
        else:
            cand, counts = out
        pp = (ctypes.c_void_p * 3)(*[_lib.ptr(p) for p in heads])
        rc = self._lib.yv3_filter_heads(ctypes.byref(self._cfg), pp, n, float(np.float32(score_thresh)), _lib.ptr(cand),
                                        self.capacity, _lib.ptr(counts), _lib.ptr(ws), ws.numel(), 3,
                                        _lib.current_stream_ptr(heads[0].device))
        _lib.check(rc, "yv3_filter_heads")
        self._last_ws = ws
        return cand, counts

    def nms(self, cand, counts, nms_thresh, out=None):
        n = cand.shape[0]
        ws = self._workspace(n, cand.device)
        keep = torch.empty((n, self.capacity), dtype=torch.uint8, device=cand.device) if out is None else out
        rc = self._lib.yv3_nms(ctypes.byref(self._cfg), _lib.ptr(cand), _lib.ptr(counts), 3, n, self.capacity,
                               max(self.class_num, 1), float(nms_thresh), _lib.ptr(keep), _lib.ptr(ws),
                               ws.numel(), _lib.current_stream_ptr(cand.device))
        _lib.check(rc, "yv3_nms")
        return keep

    def run(self, predicts, score_thresh, nms_thresh):
        cand, counts = self.filter(predicts, score_thresh)
        keep = self.nms(cand, counts, nms_thresh)
        return dict(cand=cand, counts=counts, keep=keep)

    def check_status(self):
        if self._last_ws is None:
            return
        rc = self._lib.yv3_read_status(_lib.ptr(self._last_ws), _lib.current_stream_ptr(self._last_ws.device))
        _lib.check(rc, "post-process")

    def collect(self, result, target_size, bug_compat=True):
        """apply_nms' returned rows + resize_boxes (yolov3_post_process.py:94-104, :162-172) on the device:
        (rows (N, capacity, 9) float64 [x0*w, y0*h, x1*w, y1*h, conf, prob, class, score, id], counts (N, 3) int32 rows
        per head).  ``target_size`` = (w, h) of the network input, as run.py passes it."""
        cand, counts, keep = result["cand"], result["counts"], result["keep"]
        n = cand.shape[0]
        rows = torch.empty((n, self.capacity, 9), dtype=torch.float64, device=cand.device)
        out_counts = torch.empty((n, 3), dtype=torch.int32, device=cand.device)
        rc = self._lib.yv3_collect_kept(_lib.ptr(cand), _lib.ptr(counts), _lib.ptr(keep), n, self.capacity,
                                        float(target_size[0]), float(target_size[1]), 1 if bug_compat else 0,
                                        _lib.ptr(rows), _lib.ptr(out_counts), _lib.current_stream_ptr(cand.device))
        _lib.check(rc, "yv3_collect_kept")
        return rows, out_counts

    def to_reference(self, result, image, bug_compat=True):
        """run.py's per-image structure after filter_boxes + apply_nms: 3 entries, (0,9) array or list of
        (9,) float64 rows (fp32 rows for class_num == 0 are promoted by the id column exactly as in the
        reference)."""
        counts = result["counts"][image].cpu().numpy().astype(np.int64)
        total = int(counts.sum())
        rows = result["cand"][image, :total].cpu().numpy()
        keep = result["keep"][image, :total].cpu().numpy().astype(bool)
        offsets = np.concatenate([[0], np.cumsum(counts)])
        kept_local = set()
        for h in range(3):
            kept_local.update(np.nonzero(keep[offsets[h]:offsets[h + 1]])[0].tolist())
        out = []
        for h in range(3):
            k = int(counts[h])
            if k == 0:
                out.append(np.empty((0, 9), dtype=np.float64))
                continue
            head = rows[offsets[h]:offsets[h + 1]]
            head = head.astype(np.float64) if self.class_num >= 1 else head
            full = np.concatenate([head, np.arange(k, dtype=float)[:, None]], axis=-1)
            sel = [i for i in range(k) if i in kept_local] if bug_compat else np.nonzero(
                keep[offsets[h]:offsets[h + 1]])[0].tolist()
            out.append([full[i] for i in sel])
        return out
