"""AP / mAP accumulation after NMS on the device: the evaluation the reference leaves as a TODO
pointing at Cartucho/mAP (run.py:78-79, README.md:137), over ``yv3_map_*`` (csrc/map_eval.cu).

    acc = MeanAPAccumulator(num_classes, image_size=(w, h), max_detections=1 << 20)
    for batch: res = post.run(predicts, conf_thresh, nms_thresh); acc.update(post, res, targets)
    ap, mean_ap, gt_count = acc.compute()
"""
import ctypes

import torch

from . import _lib


class MeanAPAccumulator(object):
    def __init__(self, num_classes, image_size, max_detections=1 << 20, min_overlap=0.5, pixel_inclusive=True,
                 score_col=7, device=None):
        """``num_classes`` 0 (the reference's single-class heads, configs.py:42) evaluates class-agnostically.
        ``image_size`` = (w, h) of the network input, as run.py passes it to resize_boxes."""
        self.class_agnostic = int(num_classes) == 0
        self.num_classes = max(int(num_classes), 1)
        self.image_size = (float(image_size[0]), float(image_size[1]))
        self.max_detections, self.min_overlap = int(max_detections), float(min_overlap)
        self.pixel_inclusive, self.score_col = bool(pixel_inclusive), int(score_col)
        self._lib = _lib.lib()
        size = ctypes.c_size_t(0)
        _lib.check(self._lib.yv3_map_state_bytes(self.max_detections, self.num_classes, ctypes.byref(size)), "yv3_map_state_bytes")
        _lib.require_cuda()
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else device
        self._state = torch.empty(size.value, dtype=torch.uint8, device=self.device)
        self.reset()

    def reset(self):
        with torch.cuda.device(self.device):
            _lib.check(self._lib.yv3_map_reset(_lib.ptr(self._state), self._state.numel(), self.max_detections, self.num_classes,
                                               _lib.current_stream_ptr(self.device)), "yv3_map_reset")
        self.images = 0

    def update_rows(self, kept_rows, kept_counts, targets):
        """``kept_rows`` (N, capacity, 9) float64, ``kept_counts`` (N, 3) int32 as ``HeadPostProcessor.collect`` returns them;
        ``targets`` (N, M*5) the label batch of the same images."""
        t = _lib.to_device_f32(targets, kept_rows.device)
        n, capacity = kept_rows.shape[0], kept_rows.shape[1]
        if kept_rows.dtype != torch.float64 or kept_rows.shape[2] != 9 or not kept_rows.is_contiguous():
            raise ValueError("kept_rows must be a contiguous (N, capacity, 9) float64 tensor")
        if t.dim() != 2 or t.shape[0] != n or t.shape[1] % 5:
            raise ValueError("targets must be (N, M*5) for the same N images")
        with torch.cuda.device(kept_rows.device):
            rc = self._lib.yv3_map_accumulate(_lib.ptr(kept_rows), _lib.ptr(kept_counts), n, capacity, _lib.ptr(t), t.shape[1] // 5,
                                              self.image_size[0], self.image_size[1], self.images, self.min_overlap,
                                              1 if self.pixel_inclusive else 0, self.score_col, 1 if self.class_agnostic else 0,
                                              _lib.ptr(self._state), _lib.current_stream_ptr(kept_rows.device))
        _lib.check(rc, "yv3_map_accumulate")
        self.images += n

    def update(self, post, result, targets, bug_compat=True):
        """one batch from ``HeadPostProcessor.run``'s result: the rows run.py would hold after apply_nms + resize_boxes
        (the shipped id-collision quirk included by default) go into the accumulator without leaving the device."""
        rows, kcounts = post.collect(result, self.image_size, bug_compat=bug_compat)
        self.update_rows(rows, kcounts, targets)

    def compute(self):
        """(ap (num_classes,) float64 with -1 for classes without ground truth, mAP float, gt_count (num_classes,) int32).
        Ends the evaluation: call reset() before accumulating again."""
        ap = torch.empty(self.num_classes, dtype=torch.float64, device=self.device)
        gt = torch.empty(self.num_classes, dtype=torch.int32, device=self.device)
        m2 = torch.empty(2, dtype=torch.float64, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(self._lib.yv3_map_finalize(_lib.ptr(self._state), self.num_classes, _lib.ptr(ap), _lib.ptr(gt), _lib.ptr(m2),
                                                  _lib.current_stream_ptr(self.device)), "yv3_map_finalize")
        m2h = m2.cpu()
        if float(m2h[1]) != 0.0:
            raise RuntimeError("more than max_detections=%d kept rows arrived: the AP would be truncated" % self.max_detections)
        return ap, float(m2h[0]), gt
