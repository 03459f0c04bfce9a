"""Drop-in for yolov3/yolov3_decoder.py::YOLOv3Decoder backed by yv3_decode (CUDA, sm_100a)."""
import ctypes

import torch

from . import _lib


class YOLOv3Decoder(object):
    """Same constructor and ``decode`` contract as the reference (yolov3_decoder.py:14, :62):
    ``decode(predicts)`` takes the packed head tensor (N, H/32, W/32, Ctot) and returns
    ``[(raw[..., 0:4], decoded, boxes)] * 3`` in /8, /16, /32 order with shapes (N,H,W,B,4),
    (N,H,W,B,5+class_num) in grid units, (N,H,W,B,4) corner boxes.

    Inputs may be CUDA tensors (results are CUDA tensors, asynchronous on the current stream) or
    NumPy arrays / CPU tensors (copied in through pinned memory, results returned as NumPy arrays,
    which is what run.py:65-67 converts them to anyway)."""

    def __init__(self, head_grid_sizes, class_num, anchor_boxes):
        self.head_grid_sizes = [(int(h), int(w)) for (h, w) in head_grid_sizes]
        self.class_num = int(class_num)
        self.box_len = 5 + self.class_num
        self.box_nums = [len(a) for a in anchor_boxes]
        self._cfg = _lib.make_config(self.head_grid_sizes, self.class_num, anchor_boxes)
        self.packed_channels = int(_lib.lib().yv3_packed_floats_per_image(ctypes.byref(self._cfg))) // (
            self.head_grid_sizes[2][0] * self.head_grid_sizes[2][1])

    def _check_predicts(self, predicts):
        h32, w32 = self.head_grid_sizes[2]
        if predicts.dim() != 4 or tuple(predicts.shape[1:]) != (h32, w32, self.packed_channels):
            raise ValueError("predicts must be (N, %d, %d, %d), got %s" %
                             (h32, w32, self.packed_channels, tuple(predicts.shape)))

    def decode(self, predicts, want_raw=True, want_decoded=True, want_boxes=True):
        host = _lib.is_host(predicts)
        p = _lib.to_device_f32(predicts)
        self._check_predicts(p)
        n = p.shape[0]
        outs = []
        ptrs = (ctypes.c_void_p * 9)()
        for h in range(3):
            gh, gw = self.head_grid_sizes[h]
            b = self.box_nums[h]
            raw = torch.empty((n, gh, gw, b, 4), dtype=torch.float32, device=p.device) if want_raw else None
            dec = torch.empty((n, gh, gw, b, self.box_len), dtype=torch.float32, device=p.device) if want_decoded else None
            box = torch.empty((n, gh, gw, b, 4), dtype=torch.float32, device=p.device) if want_boxes else None
            outs.append((raw, dec, box))
            ptrs[3 * h + 0] = _lib.ptr(raw)
            ptrs[3 * h + 1] = _lib.ptr(dec)
            ptrs[3 * h + 2] = _lib.ptr(box)
        rc = _lib.lib().yv3_decode(ctypes.byref(self._cfg), _lib.ptr(p), n, ptrs, _lib.current_stream_ptr(p.device))
        _lib.check(rc, "yv3_decode")
        if host:
            outs = [tuple(None if t is None else t.cpu().numpy() for t in o) for o in outs]
        return outs
