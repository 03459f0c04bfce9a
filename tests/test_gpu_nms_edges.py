"""Edge cases of the NMS kernels against the C oracle (oracle/nms_oracle.c = _apply_nms :108-159, _cal_iou :132-145 in
fp64): every path that leaves the fast fp32 decision has to give the reference's survivors index for index.

  * IoU exactly AT the threshold (strict '>', T9) and a hair above / below it: inside the +-2^-18 band the fp32 test is
    not trusted and the exact fp64 expression decides;
  * boxes so small that their area leaves the range the fp32 bounds are proved for: the whole segment goes pair by pair
    through the exact expression;
  * thresholds the fast path refuses: 0, 1e-9 and a negative one (disjoint boxes suppress each other too, :124);
  * heavy suppression (few survivors per block of 32: later blocks visit the survivors only);
  * a list the counting sort cannot bucket (one class above 256 candidates): every segment comes through the general sort,
    already sorted and without a length;
  * the same cases as ONE long single-class segment (the cluster kernel's staged kept list).
"""
import numpy as np
import pytest
import torch

from oracle import nms_c

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def yv3():
    import yolov3_tensorflow_b200 as pkg
    from yolov3_tensorflow_b200 import _lib
    _lib.lib()
    return pkg


def _post(yv3, workload):
    import bench
    spec = bench.build_spec(bench.WORKLOADS[workload])
    return yv3.HeadPostProcessor(spec.head_grid_sizes, spec.class_num, spec.anchor_boxes)


def _rows(boxes, scores, cls):
    boxes = np.asarray(boxes, dtype=np.float32)
    r = np.zeros((len(boxes), 8), dtype=np.float32)
    r[:, 0:4] = boxes
    r[:, 4] = 1.0
    r[:, 5] = scores
    r[:, 6] = cls
    r[:, 7] = scores
    return r


def _threshold_family(rng, count, cls, eps_rel):
    """pairs (A, B_k) with IoU = 0.5 * (1 + k * eps_rel), k = -2..2:  A = [0, y, 2, y + 1], B = [0, y, 1 + d, y + 1] with
    d = k * eps_rel, so inter = 1 + d and union = 2 exactly; the pairs are stacked 4 apart in y (no pair meets another)
    and everything is scaled by 2^-10: every coordinate, area and product is exact in fp32 and in fp64."""
    boxes, scores = [], []
    for i in range(count):
        y = 4.0 * i
        d = (i % 5 - 2) * eps_rel
        boxes += [[0.0, y, 2.0, y + 1.0], [0.0, y, 1.0 + d, y + 1.0]]
        scores += [0.9 - 1e-4 * i, 0.5 - 1e-4 * i]                     # A is visited first and kept; B is the one decided
    b64 = np.array(boxes, dtype=np.float64) / 1024.0
    assert np.array_equal(b64.astype(np.float32).astype(np.float64), b64)
    return _rows(b64, np.array(scores), cls)


def _crowd(rng, ln, cls, spread=0.05, scale=1.0, offset=0.0):
    cx = rng.uniform(0.3, 0.7, size=ln) + rng.normal(0, spread, size=ln)
    cy = rng.uniform(0.3, 0.7, size=ln) + rng.normal(0, spread, size=ln)
    w = rng.uniform(0.05, 0.3, size=ln)
    h = rng.uniform(0.05, 0.3, size=ln)
    b = np.stack([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2], 1) * scale + offset
    return _rows(b, rng.uniform(0.2, 1.0, size=ln), cls)


def _run(post, lists, thr):
    n = len(lists)
    cand = np.zeros((n, post.capacity, 8), dtype=np.float32)
    counts = np.zeros((n, 3), dtype=np.int32)
    for i, rows in enumerate(lists):
        k = len(rows)
        assert k <= post.capacity
        cand[i, :k] = rows
        counts[i] = [k // 2, k - k // 2, 0]
    keep = post.nms(torch.from_numpy(cand).cuda(), torch.from_numpy(counts).cuda(), thr).cpu().numpy()
    post.check_status()
    kept = []
    for i, rows in enumerate(lists):
        k = len(rows)
        got = np.nonzero(keep[i, :k])[0]
        want = np.sort(nms_c.greedy_nms(cand[i, :k].astype(np.float64), thr))
        assert np.array_equal(got, want), "list %d: survivors differ from the oracle (%d vs %d)" % (i, len(got), len(want))
        kept.append(len(got))
    return kept


def _many_class_lists(rng, general=False):
    """several classes of 33..256 candidates per image, each an edge family; general=True adds a class of 300 so that
    the counting sort gives up and every segment takes the general path"""
    lists = []
    for eps in (2.0 ** -23, 2.0 ** -20, 2.0 ** -17):
        parts = [_threshold_family(rng, 60, 3, eps), _threshold_family(rng, 100, 7, eps),
                 _crowd(rng, 200, 11), _crowd(rng, 64, 12, spread=0.01),            # heavy suppression: few survive
                 _crowd(rng, 90, 20, scale=1e-12), _crowd(rng, 40, 21, scale=1e-12, offset=0.5),   # areas ~1e-26
                 _crowd(rng, 33, 40), _crowd(rng, 256, 41), _crowd(rng, 17, 42)]
        if general:
            parts.append(_crowd(rng, 300, 50))
        rows = np.concatenate(parts, 0)
        lists.append(rows[rng.permutation(len(rows))])
    return lists


@pytest.mark.parametrize("thr", [0.5, 0.45, 1e-9, 0.0, -0.25])
def test_warp_nms_edge_cases(yv3, thr):
    post = _post(yv3, "cfg2_416_c80_b64")
    kept = _run(post, _many_class_lists(np.random.default_rng(3)), thr)
    assert all(k > 0 for k in kept)


@pytest.mark.parametrize("thr", [0.5, 0.45, 0.0])
def test_segments_through_the_general_sort(yv3, thr):
    post = _post(yv3, "cfg2_416_c80_b64")
    _run(post, _many_class_lists(np.random.default_rng(4), general=True), thr)


def test_threshold_family_is_decided_like_the_reference(yv3):
    """the family itself, at thr = 0.5: B survives for d <= 0 (IoU <= 0.5, strict '>') and is dropped for d > 0 -- whatever
    the size of d"""
    post = _post(yv3, "cfg2_416_c80_b64")
    for eps in (2.0 ** -23, 2.0 ** -20, 2.0 ** -17, 2.0 ** -10):
        rows = _threshold_family(np.random.default_rng(0), 100, 5, eps)
        kept = _run(post, [rows], 0.5)
        # per group of five pairs: A x 5 survive, B survives for k = -2, -1, 0
        assert kept[0] == 100 + 60


def _small_class_lists(rng):
    """every class at most 32 candidates: resolved inside the sort kernel, one warp per class"""
    lists = []
    for eps in (2.0 ** -23, 2.0 ** -20, 2.0 ** -17):
        parts = []
        for c in range(0, 30):
            kind = c % 5
            if kind == 0:
                parts.append(_threshold_family(rng, 5 + c % 11, c, eps))            # 10..32 candidates, IoU at / around 0.5
            elif kind == 1:
                parts.append(_crowd(rng, 3 + c, c))
            elif kind == 2:
                parts.append(_crowd(rng, 32, c, spread=0.01))                        # heavy suppression
            elif kind == 3:
                parts.append(_crowd(rng, 20, c, scale=1e-12))                        # areas ~1e-26
            else:
                parts.append(_crowd(rng, 1 + c % 3, c))
        rows = np.concatenate(parts, 0)
        lists.append(rows[rng.permutation(len(rows))])
    return lists


@pytest.mark.parametrize("thr", [0.5, 0.45, 1e-9, 0.0, -0.25])
def test_small_segments_inside_the_sort_kernel(yv3, thr):
    post = _post(yv3, "cfg2_416_c80_b64")
    kept = _run(post, _small_class_lists(np.random.default_rng(7)), thr)
    assert all(k > 0 for k in kept)


def test_small_threshold_families_are_decided_like_the_reference(yv3):
    post = _post(yv3, "cfg2_416_c80_b64")
    for eps in (2.0 ** -23, 2.0 ** -20, 2.0 ** -17, 2.0 ** -10):
        rows = np.concatenate([_threshold_family(np.random.default_rng(0), 15, c, eps) for c in range(8)], 0)
        kept = _run(post, [rows], 0.5)
        assert kept[0] == 8 * (15 + 9)                       # per class: 15 A's and the B's with k = -2, -1, 0


@pytest.mark.parametrize("thr", [0.5, 0.4, 1e-9, -0.25])
def test_long_single_class_segment_edge_cases(yv3, thr):
    post = _post(yv3, "cfg5_nms_stress_b16")
    rng = np.random.default_rng(5)
    lists = []
    for eps in (2.0 ** -23, 2.0 ** -19):
        fam = _threshold_family(rng, 400, 0, eps)
        fam[:, 0:4] *= 0.25                                                 # the families in one corner of the image
        crowd = _crowd(rng, 2500, 0, spread=0.15)
        crowd[:, 2:4] = crowd[:, 0:2] + (crowd[:, 2:4] - crowd[:, 0:2]) * 0.3
        crowd[:, 0:4] += 1.0                                                # ... the crowd elsewhere
        rows = np.concatenate([fam, crowd], 0)
        lists.append(rows[rng.permutation(len(rows))])
    tiny = _crowd(rng, 1500, 0, spread=0.15, scale=1e-12)                   # areas ~1e-26: the exact walk for everyone
    lists.append(tiny)
    _run(post, lists, thr)
