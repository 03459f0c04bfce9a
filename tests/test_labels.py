"""Label ingestion: parse -> letter-box transform -> -1 padded batch (dataset/file_util.py:22-52,97)."""
import os

import numpy as np
import pytest

from oracle import ref_labels as rl

GOLD = os.path.join(os.path.dirname(__file__), "golden", "labels_test_sample.npz")


def _load():
    z = np.load(GOLD)
    labels = [z["flat"][5 * z["offsets"][i]:5 * z["offsets"][i + 1]] for i in range(len(z["names"]))]
    return z, labels


def test_oracle_reproduces_the_fixture_and_the_parser_the_numbers():
    z, labels = _load()
    assert len(labels) == 20 and int(z["offsets"][-1]) == 51                     # the reference's demo set
    for line, name, lab in zip(z["lines"], z["names"], labels):
        n2, l2 = rl.parse_line(str(line))
        assert n2 == str(name) and np.array_equal(l2, lab)
    assert np.array_equal(rl.batch(labels, z["shapes"], (416, 416)), z["batch_416"])
    assert np.array_equal(rl.batch(labels, z["shapes"], (384, 480)), z["batch_384x480"])
    assert z["batch_416"].shape == (20, 40) and np.all(z["batch_416"][1, 5:] == -1.0)


def test_letterbox_properties():
    """A picture with the target's aspect ratio is left alone; a wide picture keeps x and squeezes y around 0.5."""
    lab = np.array([0.25, 0.75, 0.5, 0.2, 3.0], np.float32)
    assert np.array_equal(rl.letterbox(lab, (300, 300), (416, 416)), lab)
    out = rl.letterbox(lab, (200, 400), (416, 416))                               # h = w / 2
    np.testing.assert_allclose(out, [0.25, 0.75 * 0.5 + 0.25, 0.5, 0.1, 3.0], rtol=1e-6)
    assert rl.padded_batch([lab, np.zeros(0, np.float32)]).tolist()[1] == [-1.0] * 5


def test_product_parser_matches_oracle(tmp_path):
    from yolov3_tensorflow_b200.label_ingest import parse_label_line, read_label_file
    z, labels = _load()
    p = tmp_path / "label.txt"
    p.write_text("\n".join(str(l) for l in z["lines"]) + "\n\n")
    got = read_label_file(str(p))
    assert [g[0] for g in got] == [str(n) for n in z["names"]]
    for (_, g), lab in zip(got, labels):
        assert g.dtype == np.float32 and np.array_equal(g, lab)
    with pytest.raises(ValueError):
        parse_label_line("a.jpg 0.1 0.2 0.3")


@pytest.mark.gpu
@pytest.mark.parametrize("size", [(416, 416), (384, 480)])
def test_label_batch_bit_exact(size):
    import torch
    from yolov3_tensorflow_b200.label_ingest import LabelBatcher
    z, labels = _load()
    ref = rl.batch(labels, z["shapes"], size)
    got = LabelBatcher(size).batch(labels, z["shapes"])
    assert np.array_equal(got.cpu().numpy(), ref)
    # ragged extremes: an image without boxes, fixed slot count, random sizes
    rng = np.random.default_rng(3)
    labs = [rng.uniform(0.05, 0.95, size=5 * k).astype(np.float32) for k in (0, 1, 7, 3)]
    shapes = [(int(rng.integers(100, 900)), int(rng.integers(100, 900))) for _ in labs]
    got = LabelBatcher(size).batch(labs, shapes, slots=9)
    assert np.array_equal(got.cpu().numpy(), rl.batch(labs, shapes, size, slots=9))
    with pytest.raises(ValueError):
        LabelBatcher(size).batch(labs, shapes, slots=4)
    torch.cuda.synchronize()


@pytest.mark.gpu
def test_real_labels_through_the_loss():
    """BASELINE.json configs[0]: the 20-image demo labels, batch 4, through ingestion -> assignment -> loss (as shipped:
    384x480, 3-2-3 anchors, class_num 0), checked against the oracle on the same targets."""
    import torch
    import yolov3_tensorflow_b200 as yv3
    from helpers import make_predicts, make_spec, product_loss_args
    from oracle import ref_torch as rt
    from yolov3_tensorflow_b200.label_ingest import LabelBatcher
    z, labels = _load()
    spec = make_spec("shipped_384x480_c0")
    yv3.backend.set_epsilon(spec.eps)
    loss = yv3.YOLOv3Loss(rectified_coord_num=-1, **product_loss_args(spec))
    flipped = 0
    for b0 in range(0, 20, 4):
        targets = LabelBatcher((384, 480)).batch(labels[b0:b0 + 4], z["shapes"][b0:b0 + 4])
        t = targets.cpu().numpy()
        assert np.array_equal(t, rl.batch(labels[b0:b0 + 4], z["shapes"][b0:b0 + 4], (384, 480)))
        p = make_predicts(spec, 4, seed=100 + b0)
        total, grad = loss.loss_and_grad(targets, torch.from_numpy(p).cuda())
        loss.check_status()
        _, detail32, grad32 = rt.loss_and_grad(t, p, spec, dtype=torch.float32)
        _, detail64, grad64 = rt.loss_and_grad(t, p, spec, dtype=torch.float64)
        g = grad.cpu().numpy()
        # decisions (cell, anchor, head, background) are the fp32 reference's, bit for bit: same gradient support
        assert np.array_equal(g == 0, grad32 == 0)
        if not np.array_equal(grad32 == 0, grad64 == 0):
            # image 16: x = 232/480 stored as fp32, times 60 columns, is 29.0 in fp32 and 28.9999992 in fp64 -- the fp64
            # restatement puts the box one cell to the left of the reference.  Only the fp32 oracle is a valid checker.
            flipped += 1
            np.testing.assert_allclose(loss.loss_detail.cpu().numpy(), detail32, rtol=2e-5, atol=2e-6)
            gmax = np.abs(grad32).max()
            assert np.all(np.abs(g - grad32) <= 1e-5 * np.abs(grad32) + 4e-6 * gmax)
            continue
        np.testing.assert_allclose(loss.loss_detail.cpu().numpy(), detail64, rtol=1e-5, atol=1e-6)
        gmax = np.abs(grad64).max()
        # (s + c) - c at c up to 59 (480 / 8 columns) quantises the xy term in fp32: the kernel is held to
        # the fp32 reference's own distance from fp64 (x4), and element-wise to a bound that covers that quantisation
        assert np.abs(g - grad64).max() <= 4 * max(np.abs(grad32 - grad64).max(), 1e-7 * gmax)
        assert np.all(np.abs(g - grad64) <= 1e-5 * np.abs(grad64) + 4 * np.abs(grad32 - grad64).max() + 1e-6 * gmax)
    assert flipped == 1
