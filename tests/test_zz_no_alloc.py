"""yv3_nms_static (csrc/static_scratch.cu): the NMS for callers that cannot allocate -- its scratch is a file-scope device
array inside the library instead of a caller workspace.  Called through the C-ABI, several calls in a row on one stream
with nothing in between, and held to yv3_nms (with a torch-allocated workspace) on the same inputs: the keep flags of
every row below the lists' counts must be equal bit for bit (the bar of every NMS test here), and equal to the C oracle's
greedy NMS.  The calls differ in shape, class count and kernel path (device-dispatched warp NMS / cluster kernel), so a
later call finds the earlier call's keys, queues and counters in the shared array.

Collected last on purpose: nothing here touches the other entry points' state.
"""
import ctypes

import numpy as np
import pytest
import torch

from oracle import nms_c

pytestmark = pytest.mark.gpu

ERR_STATIC_SCRATCH = -10
STATIC_BYTES = 4 << 20


@pytest.fixture(scope="module")
def lib():
    from yolov3_tensorflow_b200 import _lib
    handle = _lib.lib()          # fails loudly if the CUDA library is missing
    return _lib, handle


def _rows(rng, ln, cls, spread):
    cx = rng.uniform(0.3, 0.7, size=ln) + rng.normal(0, spread, size=ln)
    cy = rng.uniform(0.3, 0.7, size=ln) + rng.normal(0, spread, size=ln)
    w = rng.uniform(0.05, 0.3, size=ln)
    h = rng.uniform(0.05, 0.3, size=ln)
    r = np.zeros((ln, 8), dtype=np.float32)
    r[:, 0], r[:, 1], r[:, 2], r[:, 3] = cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2
    r[:, 4] = rng.uniform(0.2, 1.0, size=ln)
    r[:, 5] = rng.uniform(0.2, 1.0, size=ln)
    r[:, 6] = cls
    r[:, 7] = r[:, 4] * r[:, 5]
    return r


def _many_class_case(seed, n=3, capacity=2048, num_classes=20, rest=(20, 110)):
    """segments of 1..256 candidates over 20 classes: resolved in the sort kernel and by the device-launched warp NMS"""
    rng = np.random.default_rng(seed)
    cand = np.zeros((n, capacity, 8), dtype=np.float32)
    counts = np.zeros((n, 3), dtype=np.int32)
    for i in range(n):
        lens = [1, 17, 32, 33, 64, 65, 200, 256] + list(rng.integers(rest[0], rest[1], size=num_classes - 8))
        rows = np.concatenate([_rows(rng, int(ln), c, 0.05) for c, ln in enumerate(lens)], 0)
        rows[::7, 7] = rows[0, 7]                                   # equal scores: order falls back to the row index
        rows = rows[rng.permutation(len(rows))]
        k = len(rows)
        assert k <= capacity
        cand[i, :k] = rows
        counts[i] = [k // 3, k // 3, k - 2 * (k // 3)]
    return cand, counts, num_classes, 0.45


def _single_class_case(seed, n=2, capacity=4096, k=3000):
    """one class of 3000 candidates per image: one long segment each (cluster kernel, spill list)"""
    rng = np.random.default_rng(seed)
    cand = np.zeros((n, capacity, 8), dtype=np.float32)
    counts = np.zeros((n, 3), dtype=np.int32)
    for i in range(n):
        rows = _rows(rng, k, 0, 0.15)
        rows[:, 2:4] = rows[:, 0:2] + (rows[:, 2:4] - rows[:, 0:2]) * 0.3
        cand[i, :k] = rows
        counts[i] = [k - 1000, 600, 400]
    return cand, counts, 1, 0.4


class _Case(object):
    """device buffers of one case, all allocated BEFORE any call is made"""

    def __init__(self, mod, handle, case):
        cand, counts, self.num_classes, self.thr = case
        self.cand_h, self.counts_h = cand, counts
        self.n, self.capacity = cand.shape[0], cand.shape[1]
        self.cand = torch.from_numpy(cand).cuda()
        self.counts = torch.from_numpy(counts).cuda()
        size = ctypes.c_size_t(0)
        mod.check(handle.yv3_query_postprocess_workspace(None, self.n, self.capacity, ctypes.byref(size)), "query")
        self.ws = torch.empty(size.value, dtype=torch.uint8, device="cuda")
        self.keep_ws = torch.full((self.n, self.capacity), 7, dtype=torch.uint8, device="cuda")
        self.keep_static = torch.full((self.n, self.capacity), 9, dtype=torch.uint8, device="cuda")

    def run_workspace(self, mod, handle, stream):
        mod.check(handle.yv3_nms(None, mod.ptr(self.cand), mod.ptr(self.counts), 3, self.n, self.capacity, self.num_classes,
                                 float(self.thr), mod.ptr(self.keep_ws), mod.ptr(self.ws), self.ws.numel(), stream), "yv3_nms")

    def run_static(self, mod, handle, stream):
        mod.check(handle.yv3_nms_static(mod.ptr(self.cand), mod.ptr(self.counts), 3, self.n, self.capacity, self.num_classes,
                                        float(self.thr), mod.ptr(self.keep_static), stream), "yv3_nms_static")

    def compare(self, oracle=True):
        a, b = self.keep_ws.cpu().numpy(), self.keep_static.cpu().numpy()
        for i in range(self.n):
            k = int(self.counts_h[i].sum())
            assert np.array_equal(a[i, :k], b[i, :k]), "keep flags of list %d differ from yv3_nms" % i
            assert np.all(b[i, k:] == 9), "rows beyond the count were written"
            got = np.nonzero(b[i, :k])[0]
            assert 0 < len(got) < k
            if oracle:
                assert np.array_equal(got, np.sort(nms_c.greedy_nms(self.cand_h[i, :k].astype(np.float64), self.thr)))


def test_static_scratch_nms_equals_workspace_nms_call_after_call(lib):
    mod, handle = lib
    torch.cuda.set_device(0)
    stream = mod.current_stream_ptr(torch.device("cuda", 0))
    cases = [_Case(mod, handle, _many_class_case(41)), _Case(mod, handle, _single_class_case(42)),
             _Case(mod, handle, _many_class_case(43, n=5, capacity=1024, num_classes=80, rest=(1, 6)))]
    for c in cases:
        c.run_workspace(mod, handle, stream)
    torch.cuda.synchronize()
    allocated = torch.cuda.memory_allocated()
    # calls in a row on one stream, nothing in between: each finds the previous call's contents in the shared array
    for c in cases:
        c.run_static(mod, handle, stream)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == allocated
    for c in cases:
        c.compare()
    # the same inputs again, in another order, over the leftovers of the round above: the same flags
    for c in cases:
        c.keep_static.fill_(9)
    for c in cases[::-1]:
        c.run_static(mod, handle, stream)
        c.run_static(mod, handle, stream)
    torch.cuda.synchronize()
    for c in cases:
        c.compare(oracle=False)


def test_static_scratch_limit_is_enforced(lib):
    mod, handle = lib
    cand = torch.zeros((1, 64, 8), device="cuda")
    counts = torch.zeros((1, 3), dtype=torch.int32, device="cuda")
    keep = torch.full((1, 64), 9, dtype=torch.uint8, device="cuda")
    stream = mod.current_stream_ptr(torch.device("cuda", 0))

    def need(n, capacity):          # the header's formula
        up = lambda v: (v + 255) // 256 * 256
        cap2 = 2
        while cap2 < capacity:
            cap2 *= 2
        r = n * capacity
        return up(8 * n * cap2) + up(16 * r) + up(4 * r) + up(16 * r) + up(8 * (r + r // 256 + 16)) + 256

    assert need(64, 1024) <= STATIC_BYTES < need(64, 2048)
    # beyond the array: refused before anything is launched (the pointers are far too small for these shapes)
    for n, capacity in ((64, 2048), (1, 1 << 18), (4096, 64)):
        assert need(n, capacity) > STATIC_BYTES
        rc = handle.yv3_nms_static(mod.ptr(cand), mod.ptr(counts), 3, n, capacity, 80, 0.45, mod.ptr(keep), stream)
        assert rc == ERR_STATIC_SCRATCH
    torch.cuda.synchronize()
    assert np.all(keep.cpu().numpy() == 9)
    msg = mod.strerror(ERR_STATIC_SCRATCH)
    assert "static scratch" in msg and "4 MiB" in msg
    assert handle.yv3_nms_static(None, mod.ptr(counts), 3, 1, 64, 80, 0.45, mod.ptr(keep), stream) == -1
    # the largest documented shape that fits does run (all lists empty: every count is 0)
    n, capacity = 64, 1024
    cand = torch.zeros((n, capacity, 8), device="cuda")
    counts = torch.zeros((n, 3), dtype=torch.int32, device="cuda")
    keep = torch.full((n, capacity), 9, dtype=torch.uint8, device="cuda")
    mod.check(handle.yv3_nms_static(mod.ptr(cand), mod.ptr(counts), 3, n, capacity, 80, 0.45, mod.ptr(keep), stream), "static")
    torch.cuda.synchronize()
    assert np.all(keep.cpu().numpy() == 9)
