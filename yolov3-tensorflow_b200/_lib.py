"""ctypes binding of libyv3head.so (include/yv3_head.h) and the torch glue around it.

torch is plumbing here: device memory, streams, torch.distributed.  All arithmetic of the head path
happens inside the library's CUDA kernels; there is no CPU or eager-torch fallback -- a missing
library or a missing GPU raises.
"""
import ctypes
import os

import numpy as np
import torch

from . import build as _build

YV3_MAX_BOX = 8
YV3_MAX_GT = 256
YV3_CAND_COLS = 8
YV3_MAX_CANDIDATES = 1 << 18

YV3_OK = 0
YV3_ERR_NULL = -1
YV3_ERR_SHAPE = -2
YV3_ERR_ALIGN = -4
YV3_ERR_WORKSPACE = -5
YV3_ERR_GT_RANGE = -8
YV3_ERR_NONFINITE = -9
YV3_ERR_CAPACITY = -7


class Yv3Error(RuntimeError):
    def __init__(self, code, where):
        self.code = code
        super().__init__("%s: yv3 error %d (%s)" % (where, code, strerror(code)))


class Yv3Config(ctypes.Structure):
    """struct yv3_config of include/yv3_head.h, field for field."""
    _fields_ = [
        ("grid_h", ctypes.c_int32 * 3),
        ("grid_w", ctypes.c_int32 * 3),
        ("box_num", ctypes.c_int32 * 3),
        ("class_num", ctypes.c_int32),
        ("anchors", ((ctypes.c_float * 2) * YV3_MAX_BOX) * 3),
        ("iou_thresh", ctypes.c_float),
        ("loss_weights", (ctypes.c_float * 5) * 3),
        ("rectified_loss_weight", ctypes.c_float * 3),
        ("is_focal_loss", ctypes.c_int32),
        ("focal_alpha", ctypes.c_float),
        ("focal_gamma", ctypes.c_float),
        ("is_tiou_recall", ctypes.c_int32),
        ("eps", ctypes.c_double),
    ]


_LIB = None

_P = ctypes.c_void_p
_SIGNATURES = {
    "yv3_abi_version": (ctypes.c_int, []),
    "yv3_strerror": (ctypes.c_char_p, [ctypes.c_int]),
    "yv3_config_check": (ctypes.c_int, [_P]),
    "yv3_packed_floats_per_image": (ctypes.c_int64, [_P]),
    "yv3_records_per_image": (ctypes.c_int64, [_P]),
    "yv3_query_workspace": (ctypes.c_int, [_P, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_size_t)]),
    "yv3_decode": (ctypes.c_int, [_P, _P, ctypes.c_int, ctypes.POINTER(_P), _P]),
    "yv3_label_decode": (ctypes.c_int, [_P, _P, ctypes.c_int, ctypes.c_int, ctypes.POINTER(_P), _P]),
    "yv3_label_batch": (ctypes.c_int, [_P, _P, _P, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, _P, _P]),
    "yv3_assign": (ctypes.c_int, [_P, _P, _P, ctypes.c_int, ctypes.c_int, _P, _P, _P, ctypes.c_size_t, _P]),
    "yv3_loss_fwd_bwd": (ctypes.c_int, [_P, _P, _P, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                        ctypes.c_float, _P, _P, _P, ctypes.c_size_t, ctypes.c_int, _P]),
    "yv3_loss_fwd_bwd_heads": (ctypes.c_int, [_P, ctypes.POINTER(_P), _P, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                              ctypes.c_int, ctypes.c_float, _P, ctypes.POINTER(_P), _P, ctypes.c_size_t,
                                              ctypes.c_int, _P]),
    "yv3_collect_kept": (ctypes.c_int, [_P, _P, _P, ctypes.c_int, ctypes.c_int32, ctypes.c_double, ctypes.c_double,
                                        ctypes.c_int, _P, _P, _P]),
    "yv3_read_status": (ctypes.c_int, [_P, _P]),
    "yv3_query_postprocess_workspace": (ctypes.c_int, [_P, ctypes.c_int, ctypes.c_int32,
                                                       ctypes.POINTER(ctypes.c_size_t)]),
    "yv3_filter_decoded": (ctypes.c_int, [_P, ctypes.POINTER(_P), ctypes.POINTER(_P), ctypes.c_float, _P,
                                          ctypes.c_int32, _P, _P, ctypes.c_size_t, _P]),
    "yv3_filter_packed": (ctypes.c_int, [_P, _P, ctypes.c_int, ctypes.c_float, _P, ctypes.c_int32, _P, _P,
                                         ctypes.c_size_t, ctypes.c_int, _P]),
    "yv3_filter_heads": (ctypes.c_int, [_P, ctypes.POINTER(_P), ctypes.c_int, ctypes.c_float, _P, ctypes.c_int32, _P,
                                        _P, ctypes.c_size_t, ctypes.c_int, _P]),
    "yv3_loss_filter_fused": (ctypes.c_int, [_P, _P, _P, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                             ctypes.c_float, _P, _P, _P, ctypes.c_size_t, ctypes.c_float, _P,
                                             ctypes.c_int32, _P, _P, ctypes.c_size_t, ctypes.POINTER(ctypes.c_int32),
                                             ctypes.c_int, _P]),
    "yv3_map_state_bytes": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_size_t)]),
    "yv3_map_reset": (ctypes.c_int, [_P, ctypes.c_size_t, ctypes.c_int, ctypes.c_int, _P]),
    "yv3_map_accumulate": (ctypes.c_int, [_P, _P, ctypes.c_int, ctypes.c_int32, _P, ctypes.c_int, ctypes.c_double, ctypes.c_double,
                                          ctypes.c_int, ctypes.c_double, ctypes.c_int, ctypes.c_int, ctypes.c_int, _P, _P]),
    "yv3_map_finalize": (ctypes.c_int, [_P, ctypes.c_int, _P, _P, _P, _P]),
    "yv3_peer_mailbox_bytes": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_size_t)]),
    "yv3_peer_allreduce": (ctypes.c_int, [ctypes.POINTER(_P), ctypes.c_int, ctypes.c_int, ctypes.c_int, _P, ctypes.c_int, _P, _P]),
    "yv3_peer_status": (ctypes.c_int, [_P, _P]),
    "yv3_loss_filter_fused_heads": (ctypes.c_int, [_P, ctypes.POINTER(_P), _P, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                                   ctypes.c_float, _P, ctypes.POINTER(_P), _P, ctypes.c_size_t, ctypes.c_float, _P,
                                                   ctypes.c_int32, _P, _P, ctypes.c_size_t, ctypes.POINTER(ctypes.c_int32),
                                                   ctypes.c_int, _P]),
    "yv3_nms": (ctypes.c_int, [_P, _P, _P, ctypes.c_int, ctypes.c_int, ctypes.c_int32, ctypes.c_int,
                               ctypes.c_double, _P, _P, ctypes.c_size_t, _P]),
    "yv3_nms_static": (ctypes.c_int, [_P, _P, ctypes.c_int, ctypes.c_int, ctypes.c_int32, ctypes.c_int, ctypes.c_double, _P, _P]),
}
EXPORTS = tuple(sorted(_SIGNATURES))


def library_path():
    return _build.LIB_PATH


def lib():
    """Load (building first if the in-tree .so is missing or stale and nvcc is present)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = _build.LIB_PATH
    if os.environ.get("YV3_LIB"):              # developer A/B: load this build of the library instead of the in-tree one
        path = os.environ["YV3_LIB"]
    elif not os.path.exists(path):
        path = _build.build_library()          # raises when nvcc is absent: no fallback path exists
    else:
        try:
            path = _build.build_library()      # rebuild only if the sources changed (file lock: safe under torchrun)
        except _build.NvccMissing:
            pass                                # no compiler on this box: use the shipped .so
        # a compile or link FAILURE propagates: an old .so that no longer matches the sources must not load silently
    handle = ctypes.CDLL(path)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(handle, name)             # AttributeError = header and library disagree
        fn.restype = res
        fn.argtypes = args
    if handle.yv3_abi_version() != 6:
        raise RuntimeError("libyv3head.so ABI version mismatch")
    _LIB = handle
    return _LIB


def strerror(code):
    return lib().yv3_strerror(int(code)).decode()


def check(code, where):
    if code != YV3_OK:
        raise Yv3Error(code, where)


def make_config(head_grid_sizes, class_num, anchor_boxes, iou_thresh=0.5, loss_weights=None,
                rectified_loss_weight=None, is_focal_loss=False, focal_alpha=0.25, focal_gamma=2.0,
                is_tiou_recall=False, eps=None):
    """Fill a yv3_config from the reference's constructor arguments (yolov3_loss.py:16-67,
    yolov3_decoder.py:14-45).  ``eps`` defaults to the package-wide epsilon (see backend.epsilon)."""
    from . import backend
    cfg = Yv3Config()
    if len(head_grid_sizes) != 3 or len(anchor_boxes) != 3:
        raise ValueError("head_grid_sizes and anchor_boxes need one entry per head (/8, /16, /32)")
    for h in range(3):
        gh, gw = head_grid_sizes[h]
        cfg.grid_h[h] = int(gh)
        cfg.grid_w[h] = int(gw)
        boxes = list(anchor_boxes[h])
        if not 1 <= len(boxes) <= YV3_MAX_BOX:
            raise ValueError("each head needs 1..%d anchor boxes" % YV3_MAX_BOX)
        cfg.box_num[h] = len(boxes)
        for b, (w, hh) in enumerate(boxes):
            cfg.anchors[h][b][0] = float(w)     # c_float store rounds once to fp32, like tf.constant(., float32)
            cfg.anchors[h][b][1] = float(hh)
    cfg.class_num = int(class_num)
    cfg.iou_thresh = float(iou_thresh)
    if loss_weights is None:
        loss_weights = [(1.0, 1.0, 1.0, 1.0, 1.0)] * 3
    lw = np.asarray(loss_weights, dtype=np.float64)
    if lw.shape != (3, 5):
        raise ValueError("loss_weights must be 3 rows (heads /8,/16,/32) of [xy, wh, noobj, obj, cls]")
    for h in range(3):
        for k in range(5):
            cfg.loss_weights[h][k] = float(lw[h, k])
    if rectified_loss_weight is None:
        rectified_loss_weight = [0.01, 0.01, 0.01]          # yolov3_loss.py:63-64
    elif len(rectified_loss_weight) != 3:
        # same condition and meaning as the reference's ValueError (yolov3_loss.py:65-66)
        raise ValueError("rectified_loss_weight must have length 3: the early correction-loss weights of "
                         "heads 8, 16, 32")
    for h in range(3):
        cfg.rectified_loss_weight[h] = float(rectified_loss_weight[h])
    cfg.is_focal_loss = 1 if is_focal_loss else 0
    cfg.focal_alpha = float(focal_alpha)
    cfg.focal_gamma = float(focal_gamma)
    cfg.is_tiou_recall = 1 if is_tiou_recall else 0
    cfg.eps = float(backend.epsilon() if eps is None else eps)
    rc = lib().yv3_config_check(ctypes.byref(cfg))
    if rc != YV3_OK:
        raise ValueError("invalid head configuration: " + strerror(rc))
    return cfg


# ---- torch glue -----------------------------------------------------------------------------------
def require_cuda():
    if not torch.cuda.is_available():
        raise RuntimeError("yolov3_tensorflow_b200 needs a CUDA device: the head path has no CPU fallback")


def current_stream_ptr(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def to_device_f32(x, device=None):
    """numpy / CPU tensor / CUDA tensor -> contiguous fp32 CUDA tensor.  Host inputs go through a pinned
    staging buffer (one cudaMemcpyAsync on the current stream)."""
    require_cuda()
    if isinstance(x, torch.Tensor):
        if x.is_cuda:
            t = x.detach()
            return t if (t.dtype == torch.float32 and t.is_contiguous()) else t.to(torch.float32).contiguous()
        host = x.detach().to(torch.float32).contiguous()
    else:
        host = torch.from_numpy(np.ascontiguousarray(np.asarray(x, dtype=np.float32)))
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else device
    if not host.is_pinned():
        host = host.pin_memory()
    return host.to(device, non_blocking=True)


def is_host(x):
    return not (isinstance(x, torch.Tensor) and x.is_cuda)
