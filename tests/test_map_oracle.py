"""CPU: the AP oracle (oracle/ref_map.py, the published algorithm of the tool run.py:78-79 names) on cases worked by hand."""
import numpy as np

from oracle import ref_map as rm


def test_voc_ap_all_point_interpolation():
    # three detections: tp, fp, tp over two ground-truth boxes -> rec 0.5, 0.5, 1.0; prec 1, 0.5, 2/3
    assert rm.voc_ap([0.5, 0.5, 1.0], [1.0, 0.5, 2.0 / 3.0]) == 0.5 * 1.0 + 0.5 * (2.0 / 3.0)
    assert rm.voc_ap([], []) == 0.0


def test_duplicates_are_false_positives_and_ties_keep_file_order():
    gt = [[(0, [10.0, 10.0, 50.0, 50.0])], [(0, [0.0, 0.0, 20.0, 20.0]), (1, [5.0, 5.0, 9.0, 9.0])]]
    det = [[(0, 0.9, [10.0, 10.0, 50.0, 50.0]), (0, 0.9, [11.0, 11.0, 50.0, 50.0])],       # same score: the first claims the box
           [(0, 0.8, [100.0, 100.0, 120.0, 120.0]), (0, 0.7, [0.0, 0.0, 20.0, 21.0])]]
    ap, m, cnt = rm.mean_ap(det, gt, 3)
    # class 0: order 0.9(tp) 0.9(fp, duplicate) 0.8(fp) 0.7(tp): rec .5 .5 .5 1 ; prec 1 .5 1/3 .5
    assert ap[0] == 0.5 * 1.0 + 0.5 * 0.5
    assert ap[1] == 0.0 and ap[2] == -1.0                     # class 1 has a box and no detection, class 2 nothing
    assert m == (ap[0] + 0.0) / 2 and cnt.tolist() == [2, 1, 0]


def test_gt_boxes_from_the_label_batch():
    row = np.array([0.5, 0.5, 0.25, 0.5, 3, -1, -1, -1, -1, -1], np.float32)
    assert rm.gt_boxes(row, 416.0, 416.0) == [(3, [156.0, 104.0, 260.0, 312.0])]
    assert rm.gt_boxes(row, 416.0, 416.0, class_agnostic=True)[0][0] == 0
