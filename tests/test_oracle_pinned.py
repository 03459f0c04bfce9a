"""CPU: the oracle restatements held to the REFERENCE's own outputs.

tests/golden/loss_*.npz and labels_reference.npz were produced by /root/reference/yolov3/{yolov3_loss,yolov3_decoder,
label_decoder}.py and dataset/file_util.py run unmodified under the torch-backed ``tensorflow`` stand-in
(tests/golden/make_loss_golden.py, oracle/tf_standin).  Bars:

  * fp64 twin           oracle/ref_torch.py and oracle/ref_numpy.py within 1e-12 (values, gradient relative to max|g|)
  * fp32                ref_torch bit for bit, forward (total, detail) and autodiff gradient: the same torch kernels in
                        the same order.
                        ref_numpy within 2e-6: NumPy's exp / log and pairwise summation differ from torch's by ulps --
                        those are the differing ops
  * integer decisions   valid rows, centre cells, best anchors, ownership, object and background masks: bit for bit
                        (ref_numpy; a near-tie within 1e-5 may differ for the ulp reason above and is counted)
  * closed-form backward (ref_numpy.loss_grad) within 1e-10 of the reference's autodiff gradient
"""
import os

import numpy as np
import pytest
import torch

from helpers import GOLDEN_DIR, LossFixture, loss_fixture_names
from oracle import ref_labels as rl
from oracle import ref_numpy as rn
from oracle import ref_torch as rt

NAMES = loss_fixture_names()
BIG = {"coco_416_c80", "coco_608_c80", "demo_416_c13"}          # fp64 NumPy restatement is slow there: torch only


def test_fixtures_exist():
    assert len(NAMES) >= 11


@pytest.mark.parametrize("name", NAMES)
def test_ref_torch_matches_reference_outputs(name):
    fx = LossFixture(name)
    for rect in fx.rects:
        sfx = "_rect%d" % int(rect)
        tot32, det32, g32 = rt.loss_and_grad(fx.targets, fx.predicts, fx.spec, rectified_active=rect, dtype=torch.float32)
        assert np.array_equal(det32, fx["detail32" + sfx]), "fp32 detail differs from the reference's"
        assert tot32 == fx["total32" + sfx]
        tot64, det64, g64 = rt.loss_and_grad(fx.targets, fx.predicts, fx.spec, rectified_active=rect, dtype=torch.float64)
        np.testing.assert_allclose(det64, fx["detail64" + sfx], rtol=1e-12, atol=1e-300)
        np.testing.assert_allclose(tot64, fx["total64" + sfx], rtol=1e-12)
        ref, mask, complete = fx.grad(rect, "64")
        gmax = np.abs(ref).max()
        assert np.abs(g64 - ref)[mask].max() <= 1e-12 * gmax
        if complete:
            assert not np.any(g64[~mask]) and not np.any(g32[~mask])        # exact zeros everywhere else
        ref32, _, _ = fx.grad(rect, "32")
        # fp32 autodiff with TF's registered sigmoid / softmax gradients: the same bits (NaNs of saturated logits included)
        assert np.array_equal(g32[mask], ref32[mask], equal_nan=True)


@pytest.mark.parametrize("name", NAMES)
def test_ref_numpy_matches_reference_outputs(name):
    fx = LossFixture(name)
    spec = fx.spec
    rect = fx.rects[0]
    sfx = "_rect%d" % int(rect)
    tot32, det32, decisions = rn.loss(fx.targets, fx.predicts, spec, rectified_active=rect, dtype=np.float32)
    np.testing.assert_allclose(det32, fx["detail32" + sfx], rtol=2e-6, atol=1e-7)
    ref_dec = fx.decisions()
    near = 0
    for i in range(fx.n):
        d, r = decisions[i], ref_dec[i]
        assert np.array_equal(d["valid_index"], r["valid_index"])
        rho3 = np.stack([r["heads"][h]["rho64"] for h in range(3)]) if len(r["valid_index"]) else np.zeros((3, 0))
        for h in range(3):
            dh, rh = d["heads"][h], r["heads"][h]
            assert np.array_equal(dh["rows"], rh["rows"]) and np.array_equal(dh["cols"], rh["cols"])
            diff = dh["best"] != rh["best"]
            assert np.all(dh["margin_best"][diff] <= 1e-5)
            near += int(diff.sum())
            diff = dh["own"] != rh["own"]
            assert np.all(d["margin_own"][diff] <= 1e-5)
            near += int(diff.sum())
            np.testing.assert_allclose(dh["rho"], rh["rho"], rtol=2e-6, atol=4e-6)
            for kind, key in (("obj", "object_mask"), ("bg", "background_mask")):
                m_ref = fx.mask(kind, h)[i]
                diff = dh[key] != m_ref
                if kind == "bg":
                    assert np.all(dh["margin_bg"][diff] <= 1e-5)
                    near += int(diff.sum())
                else:
                    assert not diff.any() or near > 0
            mx = fx["maxiou%d" % h][i]
            with np.errstate(invalid="ignore"):
                np.testing.assert_allclose(dh["max_iou"], mx, rtol=2e-6, atol=4e-6)   # corners carry ulp(cell index) ~ 7.6e-6 at column 75
    assert near == 0, "near-tie decisions differing between NumPy and torch exp/log: %d" % near
    if name not in BIG:
        tot64, det64, _ = rn.loss(fx.targets, fx.predicts, spec, rectified_active=rect, dtype=np.float64)
        np.testing.assert_allclose(det64, fx["detail64" + sfx], rtol=1e-12, atol=1e-300)
        g = rn.loss_grad(fx.targets, fx.predicts, spec, rectified_active=rect, dtype=np.float64)
        ref, mask, complete = fx.grad(rect, "64")
        assert np.abs(g - ref)[mask].max() <= 1e-10 * np.abs(ref).max()
        if complete:
            assert not np.any(g[~mask])


@pytest.mark.parametrize("name", NAMES)
def test_decode_and_label_decode_match_reference_outputs(name):
    fx = LossFixture(name)
    spec = fx.spec
    dec32 = rn.decode(fx.predicts, spec, np.float32)
    dec64 = rn.decode(fx.predicts, spec, np.float64)
    lab = rn.label_decode(fx.targets, spec)
    tdec = rt.decode(torch.from_numpy(fx.predicts), spec)
    for h in range(3):
        idx = fx["sample%d" % h]
        L = spec.box_len
        assert np.array_equal(dec32[h][0].reshape(-1, 4)[idx], fx["raw%d_32" % h])
        np.testing.assert_allclose(dec64[h][1].reshape(-1, L)[idx], fx["decoded%d_64" % h], rtol=1e-12)
        np.testing.assert_allclose(dec64[h][2].reshape(-1, 4)[idx], fx["boxes%d_64" % h], rtol=1e-12, atol=1e-12)
        np.testing.assert_allclose(dec32[h][1].reshape(-1, L)[idx], fx["decoded%d_32" % h], rtol=1e-6)
        # ref_torch: the same torch kernels as the stand-in -> bit for bit
        assert np.array_equal(tdec[h][1].numpy().reshape(-1, L)[idx], fx["decoded%d_32" % h])
        assert np.array_equal(tdec[h][2].numpy().reshape(-1, 4)[idx], fx["boxes%d_32" % h])
        assert np.array_equal(lab[h][0], fx["lab%d_32" % h])
        assert np.array_equal(lab[h][1], fx["labbox%d_32" % h])


def test_label_pipeline_matches_reference_outputs():
    """parse + letter-box (dataset/file_util.py:22-32, :47-52) on the 20 demo label rows, bit for bit."""
    z = np.load(os.path.join(GOLDEN_DIR, "labels_reference.npz"))
    labels = [rl.parse_line(str(line))[1] for line in z["lines"]]
    for tag, size in (("416", (416, 416)), ("384x480", (384, 480))):
        flat = np.concatenate([rl.letterbox(l, hw, size) for l, hw in zip(labels, z["shapes"])])
        assert np.array_equal(flat, z["flat_" + tag])
        assert np.array_equal(rl.batch(labels, z["shapes"], size), z["batch_" + tag])
