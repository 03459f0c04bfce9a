"""Generates tests/golden/labels_test_sample.npz from the reference's demo data (dataset/test_sample: label.txt and
the 20 pictures' sizes read with cv2), with the oracle's letter-boxed, padded batches at 416x416 (BASELINE.json
configs[0] as named) and 384x480 (as shipped, configs.py:36).  Only runnable in the build container.

    python tests/golden/make_labels_golden.py
"""
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_labels as rl  # noqa: E402

SRC = "/root/reference/dataset/test_sample"
names, labels, shapes, lines = [], [], [], []
for line in open(os.path.join(SRC, "label.txt")):
    if not line.strip():
        continue
    name, lab = rl.parse_line(line)
    img = cv2.imread(os.path.join(SRC, "images", name))
    names.append(name)
    labels.append(lab)
    shapes.append(img.shape[:2])
    lines.append(line.rstrip("\n"))
flat = np.concatenate(labels)
offsets = np.cumsum([0] + [len(l) // 5 for l in labels]).astype(np.int32)
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "labels_test_sample.npz"),
                    lines=np.array(lines), names=np.array(names), flat=flat, offsets=offsets,
                    shapes=np.array(shapes, dtype=np.int32),
                    batch_416=rl.batch(labels, shapes, (416, 416)), batch_384x480=rl.batch(labels, shapes, (384, 480)))
print(len(names), "images,", int(offsets[-1]), "boxes, longest", max(len(l) // 5 for l in labels))
