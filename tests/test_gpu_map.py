"""GPU: AP / mAP accumulated on the device (yv3_map_*) against the NumPy restatement of the tool the
reference names (oracle/ref_map.py).  Both sides do the tool's arithmetic in float64 in its operation order, so every AP
and the mean must be EQUAL, not close."""
import numpy as np
import pytest
import torch

from helpers import make_predicts, make_spec, make_targets
from oracle import ref_map as rm

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def yv3():
    import yolov3_tensorflow_b200 as pkg
    from yolov3_tensorflow_b200 import _lib
    _lib.lib()
    return pkg


def _rows_to_lists(rows, kcounts, class_agnostic, score_col=7):
    out = []
    for i in range(rows.shape[0]):
        k = int(kcounts[i].sum())
        out.append([(0 if class_agnostic else int(r[6]), float(np.float32(r[score_col])), [float(v) for v in r[:4]]) for r in rows[i, :k]])
    return out


@pytest.mark.parametrize("name,thr,inclusive", [("coco_64x96_c4", 0.05, True), ("coco_128_c3_focal_tiou", 0.1, False),
                                                ("shipped_384x480_c0", 0.6, True)])
def test_map_from_the_post_processor_matches_the_oracle(yv3, name, thr, inclusive):
    spec = make_spec(name)
    yv3.backend.set_epsilon(spec.eps)
    H, W = spec.head_grid_sizes[0][0] * 8, spec.head_grid_sizes[0][1] * 8
    post = yv3.HeadPostProcessor(spec.head_grid_sizes, spec.class_num, spec.anchor_boxes)
    acc = yv3.MeanAPAccumulator(spec.class_num, (W, H), max_detections=1 << 16, pixel_inclusive=inclusive)
    agnostic = spec.class_num == 0
    dets, gts = [], []
    for b in range(3):                                           # three batches: image order carries over
        n = 4
        p = make_predicts(spec, n, seed=40 + b)
        t = make_targets(spec, n, 6, seed=50 + b, min_valid=0, pad_in_middle=True)
        if b == 1:
            t = t.reshape(n, 6, 5)
            t[0, :, :] = -1.0                                    # an image without labels: every detection a false positive
            t = t.reshape(n, 30)
        res = post.run(torch.from_numpy(p).cuda(), thr, 0.45)
        rows, kcounts = post.collect(res, (W, H))
        acc.update_rows(rows, kcounts, torch.from_numpy(t).cuda())
        dets += _rows_to_lists(rows.cpu().numpy(), kcounts.cpu().numpy(), agnostic)
        gts += [rm.gt_boxes(t[i], float(W), float(H), agnostic) for i in range(n)]
    ap, m, cnt = acc.compute()
    ap_ref, m_ref, cnt_ref = rm.mean_ap(dets, gts, max(spec.class_num, 1), inclusive=inclusive)
    assert np.array_equal(cnt.cpu().numpy(), cnt_ref)
    assert np.array_equal(ap.cpu().numpy(), ap_ref), (ap.cpu().numpy(), ap_ref)
    assert m == m_ref
    assert sum(len(d) for d in dets) > 20


def test_map_on_crafted_detections_with_true_positives_and_ties(yv3):
    """Detections built around the labels (jittered copies, duplicates, equal scores) so that TP / FP, the duplicate rule and
    the stable tie order all matter; thousands of rows in one class exercise the global-memory sort."""
    rng = np.random.default_rng(3)
    n, m, C, cap, W, H = 16, 12, 5, 600, 416.0, 320.0
    t = np.full((n, m, 5), -1.0, np.float32)
    rows = np.zeros((n, cap, 9), np.float64)
    kc = np.zeros((n, 3), np.int32)
    for i in range(n):
        v = int(rng.integers(0, m + 1))
        xy = rng.uniform(0.1, 0.9, (v, 2))
        wh = rng.uniform(0.05, 0.3, (v, 2))
        t[i, :v, 0:2], t[i, :v, 2:4], t[i, :v, 4] = xy, wh, rng.integers(0, C, v)
        k = int(rng.integers(300, cap + 1))
        r = np.zeros((k, 9))
        for d in range(k):
            if v and rng.random() < 0.6:
                g = int(rng.integers(0, v))
                jit = rng.normal(0, 0.03, 4)
                x, y, w, h = np.float32(t[i, g, 0:4]).astype(np.float64)
                box = [(x - w / 2 + jit[0]) * W, (y - h / 2 + jit[1]) * H, (x + w / 2 + jit[2]) * W, (y + h / 2 + jit[3]) * H]
                cls = int(t[i, g, 4]) if rng.random() < 0.9 else int(rng.integers(0, C))
            else:
                x0, y0 = rng.uniform(0, W * 0.8), rng.uniform(0, H * 0.8)
                box = [x0, y0, x0 + rng.uniform(5, 100), y0 + rng.uniform(5, 100)]
                cls = int(rng.integers(0, C)) if rng.random() < 0.5 else 0
            sc = np.float32(rng.choice([0.3, 0.5, 0.7]) if rng.random() < 0.3 else rng.uniform(0.05, 1.0))
            r[d] = box + [float(sc), 1.0, cls, float(sc), d]
        rows[i, :k] = r
        kc[i] = [k // 2, k // 4, k - k // 2 - k // 4]
    t = t.reshape(n, m * 5)
    acc = yv3.MeanAPAccumulator(C, (W, H), max_detections=n * cap)
    half = n // 2
    acc.update_rows(torch.from_numpy(rows[:half]).cuda(), torch.from_numpy(kc[:half]).cuda(), torch.from_numpy(t[:half]).cuda())
    acc.update_rows(torch.from_numpy(rows[half:]).cuda(), torch.from_numpy(kc[half:]).cuda(), torch.from_numpy(t[half:]).cuda())
    ap, mean, cnt = acc.compute()
    ap_ref, m_ref, cnt_ref = rm.mean_ap(_rows_to_lists(rows, kc, False), [rm.gt_boxes(t[i], W, H) for i in range(n)], C)
    assert np.array_equal(cnt.cpu().numpy(), cnt_ref)
    assert np.array_equal(ap.cpu().numpy(), ap_ref), (ap.cpu().numpy(), ap_ref)
    assert mean == m_ref and 0.05 < mean < 0.95
    # too small an accumulator is reported, not silently truncated
    small = yv3.MeanAPAccumulator(C, (W, H), max_detections=100)
    small.update_rows(torch.from_numpy(rows[:2]).cuda(), torch.from_numpy(kc[:2]).cuda(), torch.from_numpy(t[:2]).cuda())
    with pytest.raises(RuntimeError):
        small.compute()
