"""Host-side readers of the (6, 3) loss detail for the reference's two observers.

The reference keeps six tf.Variables of shape (3,) that YOLOv3Loss.loss assigns every step (yolov3_loss.py:71-79,
:115-134); utils/logger_callback.py:114-128 evaluates them for the console line and utils/board_callback.py:52-70
registers them as TensorBoard scalars.  Here the fused call leaves the same 18 numbers in ``loss.loss_detail`` (rows
xy, wh, noobj, obj, cls, rectified x heads /8, /16, /32); these helpers read them with ONE device-to-host copy and
hand them over under the reference's names and in its order.
"""
import numpy as np

# row of loss_detail for each name the observers use
_ROWS = {"xy_loss": 0, "wh_loss": 1, "noobj_iou_loss": 2, "obj_iou_loss": 3, "cls_loss": 4, "rectified_loss": 5}
# the order of utils/logger_callback.py:121-127
_CONSOLE_ORDER = ("rectified_loss", "xy_loss", "wh_loss", "noobj_iou_loss", "obj_iou_loss", "cls_loss")
# board_callback.py:62-68 names the class term "class_loss"
_BOARD_NAMES = (("rectified_loss", "rectified_loss"), ("xy_loss", "xy_loss"), ("wh_loss", "wh_loss"),
                ("noobj_iou_loss", "noobj_iou_loss"), ("obj_iou_loss", "obj_iou_loss"), ("class_loss", "cls_loss"))


def _detail_array(loss_or_detail):
    d = getattr(loss_or_detail, "loss_detail", loss_or_detail)
    if d is None:
        raise ValueError("no loss detail yet: call loss() / loss_and_grad() first")
    if hasattr(d, "detach"):
        d = d.detach().cpu().numpy()
    d = np.asarray(d, dtype=np.float32)
    if d.shape != (6, 3):
        raise ValueError("loss detail must be (6, 3), got %s" % (d.shape,))
    return d


def detail_log_values(loss_or_detail):
    """The (key, value) pairs utils/logger_callback.py:120-128 appends, in its order:
    ('head', '/8:\\n'), ('rectified_loss', v), ('xy_loss', v), ... ('cls_loss', v), (' ', '\\n'), then /16, /32."""
    d = _detail_array(loss_or_detail)
    out = []
    for i, head in enumerate(("/8:\n", "/16:\n", "/32:\n")):
        out.append(("head", head))
        for key in _CONSOLE_ORDER:
            out.append((key, float(d[_ROWS[key], i])))
        out.append((" ", "\n"))
    return out


def format_detail(loss_or_detail):
    """The console text of logger_callback.py:130-136: strings verbatim, numbers as ' - %s: %.4f' when > 1e-3 and as
    ' - %s: %.4e' otherwise (nothing is hidden; only the format switches)."""
    info = ""
    for key, value in detail_log_values(loss_or_detail):
        if isinstance(value, str):
            info += " - %s: %s" % (key, value)
        elif value > 1e-3:
            info += " - %s: %.4f" % (key, value)
        else:
            info += " - %s: %.4e" % (key, value)
    return info


def detail_scalars(loss_or_detail):
    """The 18 TensorBoard scalars of board_callback.py:60-68: {'head_8_rectified_loss': v, 'head_8_xy_loss': v, ...}."""
    d = _detail_array(loss_or_detail)
    out = {}
    for i, name in enumerate(("head_8", "head_16", "head_32")):
        for board, key in _BOARD_NAMES:
            out["%s_%s" % (name, board)] = float(d[_ROWS[key], i])
    return out
