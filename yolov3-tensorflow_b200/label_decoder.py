"""Drop-in for yolov3/label_decoder.py::LabelDecoder backed by yv3_label_decode."""
import ctypes

import torch

from . import _lib

# LabelDecoder never looks at anchors or classes; the C config needs some, so a neutral prior is used.
_NEUTRAL_ANCHORS = [[(1.0, 1.0)]] * 3


class LabelDecoder(object):
    """``decode(targets)``: (N, M*5) rows ``[x, y, w, h, cls] * M`` normalised to the input (padding -1,
    dataset/file_util.py:97) -> ``[(targets_h (N,M,5), boxes_h (N,M,4))] * 3`` in each head's grid units
    (label_decoder.py:26-60)."""

    def __init__(self, head_grid_sizes):
        self.head_grid_sizes = [(int(h), int(w)) for (h, w) in head_grid_sizes]
        self._cfg = _lib.make_config(self.head_grid_sizes, 0, _NEUTRAL_ANCHORS)

    def decode(self, targets):
        host = _lib.is_host(targets)
        t = _lib.to_device_f32(targets)
        if t.dim() != 2 or t.shape[1] % 5 != 0 or t.shape[1] == 0:
            raise ValueError("targets must be (N, M*5)")
        n, m = t.shape[0], t.shape[1] // 5
        outs = []
        ptrs = (ctypes.c_void_p * 6)()
        for h in range(3):
            lab = torch.empty((n, m, 5), dtype=torch.float32, device=t.device)
            box = torch.empty((n, m, 4), dtype=torch.float32, device=t.device)
            outs.append((lab, box))
            ptrs[2 * h] = _lib.ptr(lab)
            ptrs[2 * h + 1] = _lib.ptr(box)
        rc = _lib.lib().yv3_label_decode(ctypes.byref(self._cfg), _lib.ptr(t), n, m, ptrs,
                                         _lib.current_stream_ptr(t.device))
        _lib.check(rc, "yv3_label_decode")
        if host:
            outs = [tuple(x.cpu().numpy() for x in o) for o in outs]
        return outs
