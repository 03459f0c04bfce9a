"""Label ingestion for the head path: the part of dataset/file_util.py that produces ``targets``.

    parse_label_line / read_label_file   FileUtil._parse_string_line (file_util.py:22-32): ``image_name x y w h cls ...``
    LabelBatcher.batch                    the letter-box label transform of _parse_image_with_label (:47-52) and the
                                          -1 padded batch of get_dataset (:97), on the GPU (yv3_label_batch)

Image decoding, resizing and augmentation stay outside the head path; the source image sizes
(``tf.shape(image)[0:2]``) come from whoever decodes the pictures.
"""
import ctypes

import numpy as np
import torch

from . import _lib


def parse_label_line(line):
    """``image_name l0 l1 ...`` -> (image_name, float32 array of 5*k numbers); split on single spaces like
    tf.string_split(delimiter=' ') (empty fields are dropped), numbers through float32 like tf.string_to_number."""
    fields = [f for f in line.rstrip("\r\n").split(" ") if f != ""]
    if not fields:
        raise ValueError("empty label line")
    values = np.array([np.float32(v) for v in fields[1:]], dtype=np.float32)
    if values.size % 5 != 0:
        raise ValueError("label line of %s: %d numbers, not a multiple of 5" % (fields[0], values.size))
    return fields[0], values


def read_label_file(path):
    """[(image_name, labels float32 (k*5,))] for every non-empty line of a label.txt."""
    out = []
    with open(path, "r") as f:
        for line in f:
            if line.strip():
                out.append(parse_label_line(line))
    return out


class LabelBatcher(object):
    """``batch(labels, image_shapes)`` -> (N, M*5) fp32 CUDA tensor, the ``targets`` of YOLOv3Loss.loss.

    ``labels``: per image a float32 array of k*5 numbers normalised to the SOURCE image; ``image_shapes``: per
    image (height, width) of the source picture; ``image_size`` (H, W): the network input (configs.py:36).
    M = the largest k of the batch (padded_batch pads to the longest element) unless ``slots`` fixes it."""

    def __init__(self, image_size, device=None):
        self.image_size = (int(image_size[0]), int(image_size[1]))
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)

    def batch(self, labels, image_shapes, slots=None):
        n = len(labels)
        if n == 0 or len(image_shapes) != n:
            raise ValueError("labels and image_shapes must be non-empty lists of the same length")
        rows = [np.asarray(l, dtype=np.float32).reshape(-1, 5) for l in labels]
        counts = [r.shape[0] for r in rows]
        m = max(max(counts), 1) if slots is None else int(slots)
        if max(counts) > m:
            raise ValueError("an image has %d label rows, more than the %d slots" % (max(counts), m))
        offsets = np.zeros(n + 1, dtype=np.int32)
        offsets[1:] = np.cumsum(counts)
        flat = np.concatenate(rows, axis=0) if offsets[-1] else np.zeros((1, 5), np.float32)
        hw = np.asarray(image_shapes, dtype=np.int32).reshape(n, 2)
        d_flat = torch.from_numpy(np.ascontiguousarray(flat)).to(self.device)
        d_off = torch.from_numpy(offsets).to(self.device)
        d_hw = torch.from_numpy(np.ascontiguousarray(hw)).to(self.device)
        targets = torch.empty((n, m * 5), dtype=torch.float32, device=self.device)
        rc = _lib.lib().yv3_label_batch(_lib.ptr(d_flat), _lib.ptr(d_off), _lib.ptr(d_hw), n, m, self.image_size[0],
                                        self.image_size[1], _lib.ptr(targets), _lib.current_stream_ptr(self.device))
        _lib.check(rc, "yv3_label_batch")
        return targets
