"""CPU: the C-ABI library loads, exports every function include/yv3_head.h declares, and its host-side
argument checks behave (no kernel is launched here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "yv3_head.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(yv3_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def lib():
    from yolov3_tensorflow_b200 import _lib
    return _lib


def test_every_declared_symbol_is_exported_and_bound(lib):
    names = _declared()
    assert len(names) >= 14
    handle = lib.lib()
    for n in names:
        assert hasattr(handle, n), "libyv3head.so does not export " + n
    assert sorted(lib.EXPORTS) == names, "the ctypes binding and the header disagree"


def test_config_struct_matches_header(lib):
    # 3*3 int32 + class_num, anchors 3*8*2 fp32, iou, 15 weights, 3 rect, focal(int, 2 fp32), tiou, pad, eps double
    assert ctypes.sizeof(lib.Yv3Config) == 336


def test_config_check_and_geometry(lib):
    from oracle.head_spec import SHIPPED_ANCHORS, coco_anchors
    cfg = lib.make_config([(48, 60), (24, 30), (12, 15)], 0, SHIPPED_ANCHORS, 0.8)
    h = lib.lib()
    assert h.yv3_records_per_image(ctypes.byref(cfg)) == 10620            # as shipped
    assert h.yv3_packed_floats_per_image(ctypes.byref(cfg)) == 10620 * 5
    cfg = lib.make_config([(52, 52), (26, 26), (13, 13)], 80, coco_anchors(), 0.8)
    assert h.yv3_records_per_image(ctypes.byref(cfg)) == 10647
    assert h.yv3_packed_floats_per_image(ctypes.byref(cfg)) == 13 * 13 * 5355
    size = ctypes.c_size_t(0)
    assert h.yv3_query_workspace(ctypes.byref(cfg), 64, 16, ctypes.byref(size)) == 0 and size.value > 0
    assert h.yv3_query_workspace(ctypes.byref(cfg), 64, 1000, ctypes.byref(size)) == -2       # m > YV3_MAX_GT
    assert h.yv3_query_postprocess_workspace(ctypes.byref(cfg), 4, 10647, ctypes.byref(size)) == 0
    assert h.yv3_query_postprocess_workspace(None, 4, 1 << 20, ctypes.byref(size)) == -7
    with pytest.raises(ValueError):
        lib.make_config([(52, 52), (26, 26), (12, 12)], 80, coco_anchors())     # not /8,/16,/32 of one input
    with pytest.raises(ValueError):
        lib.make_config([(52, 52), (26, 26), (13, 13)], 80, [[(0.1, 0.1)] * 9] * 3)   # > YV3_MAX_BOX
    with pytest.raises(ValueError):
        lib.make_config([(52, 52), (26, 26), (13, 13)], 80, coco_anchors(), float("nan"))
    assert b"NULL" in h.yv3_strerror(-1)
    # NULL pointers are refused before anything touches the device
    assert h.yv3_loss_fwd_bwd(ctypes.byref(cfg), None, None, 1, 1, 1, 0, 1.0, None, None, None, 0, 3, None) == -1


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "yolov3-tensorflow_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("# oracle", ""), f + " mentions the oracle"


def test_configs_surface():
    from yolov3_tensorflow_b200.configs import FLAGS
    assert [tuple(g) for g in FLAGS.head_grid_sizes] == [(48, 60), (24, 30), (12, 15)]
    assert FLAGS.box_len == 5 and FLAGS.rectified_coord_num == 1464 and FLAGS.nms_thresh == 0.4
