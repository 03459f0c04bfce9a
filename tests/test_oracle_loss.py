"""CPU checks of the loss / decode oracle itself (no GPU, no product code).

The reference's TF graph cannot run here, so the two independent restatements
(oracle/ref_numpy.py, oracle/ref_torch.py) are checked against each other, and the closed-form backward
of SURVEY Appendix A against torch autograd of the restated graph in fp64.
"""
import numpy as np
import pytest
import torch

from oracle import ref_numpy as rn
from oracle import ref_torch as rt

from helpers import make_predicts, make_spec, make_targets

SMALL = ["coco_64x96_c4", "coco_128_c3_focal_tiou", "tiny_32_c2", "single_160_c1_focal"]


@pytest.mark.parametrize("name", SMALL)
def test_numpy_and_torch_restatements_agree(name):
    spec = make_spec(name)
    p = make_predicts(spec, 3, seed=1)
    t = make_targets(spec, 3, 6, seed=2, min_valid=0, pad_in_middle=True)
    total_np, detail_np, _ = rn.loss(t, p, spec, rectified_active=True, dtype=np.float64)
    total_t, detail_t = rt.loss(torch.from_numpy(t).double(), torch.from_numpy(p).double(), spec, rectified_active=True)
    np.testing.assert_allclose(detail_np, detail_t.numpy(), rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(total_np, total_t.item(), rtol=1e-12)
    # fp32 restatements agree to fp32 noise
    total32, detail32, _ = rn.loss(t, p, spec, rectified_active=True, dtype=np.float32)
    np.testing.assert_allclose(detail32, detail_np, rtol=2e-5, atol=1e-6)


@pytest.mark.parametrize("name", SMALL)
def test_analytic_backward_matches_autograd_fp64(name):
    spec = make_spec(name)
    p = make_predicts(spec, 2, seed=3)
    t = make_targets(spec, 2, 5, seed=4, min_valid=1)
    # two labels on the same cell/anchor (rule T3) and a class id outside [0, C)
    t = t.reshape(2, 5, 5)
    t[0, 1] = t[0, 0]
    t[1, 0, 4] = spec.class_num + 3
    t = t.reshape(2, 25)
    g = rn.loss_grad(t, p, spec, rectified_active=True, upstream=1.0, dtype=np.float64)
    _, _, g_auto = rt.loss_and_grad(t, p, spec, rectified_active=True, dtype=torch.float64)
    scale = np.abs(g_auto).max()
    assert np.abs(g - g_auto).max() <= 1e-10 * max(scale, 1.0)


def test_decode_pack_roundtrip():
    spec = make_spec("coco_64x96_c4")
    p = make_predicts(spec, 2, seed=5)
    heads = rn.unpack(p, spec)
    back = rn.pack_heads(*[h.reshape(h.shape[0], h.shape[1], h.shape[2], -1) for h in heads])
    assert np.array_equal(back, p)
    # Appendix C offset formula
    rng = np.random.default_rng(0)
    flat = p.reshape(-1)
    for _ in range(200):
        h = int(rng.integers(0, 3))
        H, W = spec.head_grid_sizes[h]
        n, y, x = int(rng.integers(0, 2)), int(rng.integers(0, H)), int(rng.integers(0, W))
        b, ch = int(rng.integers(0, spec.box_num[h])), int(rng.integers(0, spec.box_len))
        assert flat[rn.packed_offset(spec, n, h, y, x, b, ch)] == heads[h][n, y, x, b, ch]


def test_empty_image_is_all_background():
    spec = make_spec("tiny_32_c2")
    p = make_predicts(spec, 1, seed=6)
    t = np.full((1, 10), -1.0, np.float32)
    _, detail, dec = rn.loss(t, p, spec)
    assert detail[0].sum() == 0 and detail[1].sum() == 0 and detail[3].sum() == 0 and detail[4].sum() == 0
    assert all(d["background_mask"].all() for d in dec[0]["heads"])
