"""Host-side observers of the (6,3) loss detail: names, order and formatting of
utils/logger_callback.py:114-134 and utils/board_callback.py:52-70."""
import numpy as np
import pytest


def _detail():
    # rows xy, wh, noobj, obj, cls, rectified x heads /8, /16, /32
    return np.array([[0.5, 0.25, 0.0005], [1.5, 2.5, 3.5], [10.0, 20.0, 30.0], [0.1, 0.2, 0.3], [4.0, 5.0, 6.0],
                     [0.0, 0.0, 0.0]], dtype=np.float32)


def test_console_values_follow_the_reference_order():
    from yolov3_tensorflow_b200 import detail_log_values
    v = detail_log_values(_detail())
    assert [k for k, _ in v[:8]] == ["head", "rectified_loss", "xy_loss", "wh_loss", "noobj_iou_loss", "obj_iou_loss",
                                    "cls_loss", " "]
    assert v[0][1] == "/8:\n" and v[8][1] == "/16:\n" and v[16][1] == "/32:\n"
    assert v[2] == ("xy_loss", 0.5) and v[10] == ("xy_loss", 0.25)
    assert v[22] == ("cls_loss", 6.0)
    assert len(v) == 24


def test_console_text_switches_format_for_small_numbers_like_the_reference():
    from yolov3_tensorflow_b200 import format_detail
    text = format_detail(_detail())
    # logger_callback.py:130-136: > 1e-3 as %.4f, everything else as %.4e -- nothing is dropped
    assert text.startswith(" - head: /8:\n - rectified_loss: 0.0000e+00 - xy_loss: 0.5000 - wh_loss: 1.5000"
                           " - noobj_iou_loss: 10.0000")
    assert " - head: /32:\n - rectified_loss: 0.0000e+00 - xy_loss: 5.0000e-04 - wh_loss: 3.5000" in text
    assert text.count("rectified_loss") == 3


def test_board_scalars_names():
    from yolov3_tensorflow_b200 import detail_scalars
    s = detail_scalars(_detail())
    assert len(s) == 18
    assert s["head_8_xy_loss"] == 0.5 and s["head_32_class_loss"] == 6.0 and s["head_16_noobj_iou_loss"] == 20.0
    assert set(k.split("_", 2)[2] for k in s) == {"rectified_loss", "xy_loss", "wh_loss", "noobj_iou_loss", "obj_iou_loss",
                                                  "class_loss"}


def test_rejects_missing_or_misshapen_detail():
    from yolov3_tensorflow_b200 import detail_scalars

    class Fresh(object):
        loss_detail = None
    with pytest.raises(ValueError):
        detail_scalars(Fresh())
    with pytest.raises(ValueError):
        detail_scalars(np.zeros((3, 6), dtype=np.float32))
