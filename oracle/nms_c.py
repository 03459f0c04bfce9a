"""TEST INFRASTRUCTURE ONLY (oracle): ctypes loader for oracle/nms_oracle.c (built by oracle/Makefile)."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "libnms_oracle.so")


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])
    return _SO


def _lib():
    if not os.path.exists(_SO):
        build()
    lib = ctypes.CDLL(_SO)
    lib.yv3o_greedy_nms.restype = ctypes.c_int64
    lib.yv3o_greedy_nms.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_double, ctypes.c_void_p]
    return lib


def greedy_nms(cands, nms_thresh):
    cands = np.ascontiguousarray(np.asarray(cands, dtype=np.float64)[:, :8])
    keep = np.empty((max(len(cands), 1),), dtype=np.int64)
    n = _lib().yv3o_greedy_nms(cands.ctypes.data, len(cands), float(nms_thresh), keep.ctypes.data)
    return keep[:n].copy()
