"""TEST INFRASTRUCTURE ONLY (oracle). Not imported by the product package.

NumPy restatement (fp32, one rounding per operator) of the label part of the reference's input pipeline:
  parse      dataset/file_util.py:22-32   tf.string_split(' ') + tf.string_to_number(float32)
  letter-box dataset/file_util.py:47-52   src_size_hw = shape[0:2] / image_size ; ratio = src_size_hw[::-1] / max ;
                                          xy * ratio + (1 - ratio) / 2 ; wh * ratio
  batch      dataset/file_util.py:97      padded_batch(..., padding_values=-1.) pads to the longest element

Parity status: **parity unpinned** -- these are tf.data / TensorFlow ops and TensorFlow cannot be imported here; the
reference has no test or fixture for them.  tests/golden/labels_test_sample.npz holds the reference's own label.txt
numbers and picture sizes (dataset/test_sample) with THIS restatement's outputs.
"""
import numpy as np

F = np.float32


def parse_line(line):
    fields = [f for f in line.rstrip("\r\n").split(" ") if f != ""]
    return fields[0], np.array([F(v) for v in fields[1:]], dtype=F)


def letterbox(label, image_hw, image_size):
    """label (k*5,) fp32, image_hw (h, w) of the source picture, image_size (H, W) -> (k*5,) fp32  (:47-52)."""
    lab = np.asarray(label, dtype=F).reshape(-1, 5)
    src = np.array([F(image_hw[0]), F(image_hw[1])], dtype=F) / np.array([F(image_size[0]), F(image_size[1])], dtype=F)
    ratio = (src[::-1] / src.max()).astype(F)                         # (w ratio, h ratio)
    xy = (lab[:, 0:2] * ratio).astype(F) + ((F(1) - ratio).astype(F) / F(2.0)).astype(F)
    wh = (lab[:, 2:4] * ratio).astype(F)
    return np.concatenate([xy.astype(F), wh, lab[:, 4:]], axis=-1).astype(F).reshape(-1)


def padded_batch(labels, slots=None):
    """list of (k_i*5,) -> (N, M*5) with -1 padding, M = longest (or `slots`)  (:97)."""
    m = max(max(len(l) // 5 for l in labels), 1) if slots is None else slots
    out = np.full((len(labels), m * 5), -1.0, dtype=F)
    for i, l in enumerate(labels):
        out[i, :len(l)] = l
    return out


def batch(labels, image_shapes, image_size, slots=None):
    return padded_batch([letterbox(l, hw, image_size) for l, hw in zip(labels, image_shapes)], slots)
