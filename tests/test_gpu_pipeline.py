"""The three-stream schedule bench.py times (yolov3_tensorflow_b200.pipeline.FusedPipeline) is checked for
correctness: K DISTINCT batches go through it -- eagerly and as a CUDA-graph replay, at every depth -- and every step's
loss row, gradient, candidate rows, counts and keep flags must equal, bit for bit, what the same calls give when they
are issued one after the other on one stream with one workspace.  A missing dependency edge between the streams (sweep
k + depth overwriting the staging rows gather k still reads, the assignment of step k + depth overwriting the loss
workspace the fold of step k still reads, ...) shows up as a difference because no two steps see the same data.

The serial schedule itself is held to the reference's outputs in tests/test_gpu_reference_fixtures.py.
"""
import numpy as np
import pytest
import torch

from helpers import make_predicts, make_spec, make_targets, product_loss_args

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def yv3():
    import yolov3_tensorflow_b200 as pkg
    from yolov3_tensorflow_b200 import _lib
    _lib.lib()
    return pkg


def _batches(spec, k, n, m, seed):
    ps = [torch.from_numpy(make_predicts(spec, n, seed=seed + 10 * i)).cuda() for i in range(k)]
    ts = [torch.from_numpy(make_targets(spec, n, m, seed=seed + 10 * i + 1, min_valid=0, pad_in_middle=True)).cuda()
          for i in range(k)]
    return ts, ps


def _buffers(pipe, ps, k, n):
    dev = ps[0].device
    return dict(grads=[torch.full_like(p, float("nan")) for p in ps],
                rows=torch.full((k, 20), float("nan"), device=dev),
                cand=[torch.zeros((n, pipe.capacity, 8), device=dev) for _ in range(k)],
                counts=[torch.full((n, 3), -1, dtype=torch.int32, device=dev) for _ in range(k)],
                keep=[torch.full((n, pipe.capacity), 7, dtype=torch.uint8, device=dev) for _ in range(k)])


def _same(a, b, k, n):
    assert torch.equal(a["rows"], b["rows"]), "loss rows differ"
    for i in range(k):
        assert torch.equal(a["grads"][i], b["grads"][i]), "gradient of step %d differs" % i
        assert torch.equal(a["counts"][i], b["counts"][i]), "counts of step %d differ" % i
        tot = a["counts"][i].sum(dim=1).cpu().numpy()
        for j in range(n):
            t = int(tot[j])
            assert torch.equal(a["cand"][i][j, :t], b["cand"][i][j, :t]), "candidates of step %d image %d differ" % (i, j)
            assert torch.equal(a["keep"][i][j, :t], b["keep"][i][j, :t]), "keep flags of step %d image %d differ" % (i, j)


@pytest.mark.parametrize("name,n,thr", [("coco_64x96_c4", 5, 0.05), ("coco_416_c80", 3, 0.1), ("single_160_c1_focal", 4, 0.3)])
@pytest.mark.parametrize("depth", [1, 2, 3, 4])
def test_pipelined_steps_equal_serial_steps(yv3, name, n, thr, depth):
    spec = make_spec(name)
    yv3.backend.set_epsilon(spec.eps)
    k, m = 9, 6
    ts, ps = _batches(spec, k, n, m, seed=100)
    # the correction window closes inside the run: steps 0..3 have the term, 4.. do not (yolov3_loss.py:125)
    rcn = 3 * n
    ref = yv3.FusedPipeline(product_loss_args(spec), thr, 0.45, depth=1, rectified_coord_num=rcn)
    want = _buffers(ref, ps, k, n)
    ref.run(ts, ps, serial=True, **want)
    torch.cuda.synchronize()
    ref.check_status()
    assert ref.current_num == 4 * n and ref.last_fused
    assert float(want["rows"][3, 15:18].abs().sum()) > 0 and float(want["rows"][4, 15:18].abs().sum()) == 0

    pipe = yv3.FusedPipeline(product_loss_args(spec), thr, 0.45, depth=depth, rectified_coord_num=rcn)
    got = _buffers(pipe, ps, k, n)
    pipe.run(ts, ps, **got)
    torch.cuda.synchronize()
    pipe.check_status()
    assert pipe.current_num == 4 * n and pipe.state_dict() == {"current_num": 4 * n}
    _same(want, got, k, n)
    # a second run continues the counter: every step now without the correction term
    got2 = _buffers(pipe, ps, k, n)
    pipe.run(ts, ps, **got2)
    ref.run(ts, ps, serial=True, **want)
    torch.cuda.synchronize()
    _same(want, got2, k, n)


@pytest.mark.parametrize("depth", [2, 3])
def test_graph_replay_equals_serial_steps(yv3, depth):
    spec = make_spec("coco_96x64_c13")
    yv3.backend.set_epsilon(spec.eps)
    k, n, m = 8, 4, 5
    ts, ps = _batches(spec, k, n, m, seed=300)
    ref = yv3.FusedPipeline(product_loss_args(spec), 0.05, 0.4, depth=1, rectified_coord_num=-1)
    want = _buffers(ref, ps, k, n)
    ref.run(ts, ps, serial=True, **want)
    pipe = yv3.FusedPipeline(product_loss_args(spec), 0.05, 0.4, depth=depth, rectified_coord_num=-1)
    got = _buffers(pipe, ps, k, n)
    graph = pipe.capture(ts, ps, **got)
    for _ in range(2):
        for key in ("grads", "cand", "counts", "keep"):              # wipe: the replay must produce everything again
            for t in got[key]:
                t.fill_(0)
        got["rows"].fill_(0)
        graph.replay()
        torch.cuda.synchronize()
        _same(want, got, k, n)
    # new data in the same buffers, same graph
    ts2, ps2 = _batches(spec, k, n, m, seed=900)
    for i in range(k):
        ps[i].copy_(ps2[i])
        ts[i].copy_(ts2[i])
    ref.run(ts, ps, serial=True, **want)
    graph.replay()
    torch.cuda.synchronize()
    _same(want, got, k, n)


def test_shared_output_buffers_like_the_benchmark(yv3):
    """bench.py hands ONE gradient / cand / counts / keep buffer to all steps: the last step's results stay."""
    spec = make_spec("coco_64_c20")
    yv3.backend.set_epsilon(spec.eps)
    k, n, m = 7, 4, 6
    ts, ps = _batches(spec, k, n, m, seed=500)
    ref = yv3.FusedPipeline(product_loss_args(spec), 0.05, 0.45, depth=1, rectified_coord_num=-1)
    want = _buffers(ref, ps, k, n)
    ref.run(ts, ps, serial=True, **want)
    pipe = yv3.FusedPipeline(product_loss_args(spec), 0.05, 0.45, depth=3, rectified_coord_num=-1)
    dev = ps[0].device
    grad = torch.empty_like(ps[0])
    rows = torch.zeros((k, 20), device=dev)
    cand = torch.zeros((n, pipe.capacity, 8), device=dev)
    counts = torch.zeros((n, 3), dtype=torch.int32, device=dev)
    keep = torch.zeros((n, pipe.capacity), dtype=torch.uint8, device=dev)
    pipe.run(ts, ps, grad, rows, cand, counts, keep)
    torch.cuda.synchronize()
    assert torch.equal(rows, want["rows"]) and torch.equal(grad, want["grads"][-1])
    assert torch.equal(counts, want["counts"][-1])
    for j in range(n):
        t = int(counts[j].sum())
        assert torch.equal(cand[j, :t], want["cand"][-1][j, :t]) and torch.equal(keep[j, :t], want["keep"][-1][j, :t])


def test_phase_calls_on_one_instance_latch_the_rectified_switch(yv3):
    """phases 1 / 2 / 8 / 4 of one step as separate calls on ONE YOLOv3Loss re-evaluated the tf.cond per call;
    the step's first call now latches it.  Window of one step: step 0 active, step 1 not -- the sweep and the fold of
    each step must agree."""
    spec = make_spec("tiny_32_c2")
    yv3.backend.set_epsilon(spec.eps)
    n, m = 3, 4
    ts, ps = _batches(spec, 2, n, m, seed=700)
    post = yv3.HeadPostProcessor(spec.head_grid_sizes, spec.class_num, spec.anchor_boxes)
    whole = yv3.YOLOv3Loss(rectified_coord_num=0, **product_loss_args(spec))
    split = yv3.YOLOv3Loss(rectified_coord_num=0, **product_loss_args(spec))
    for k in range(2):
        tot_w, grad_w, _, _ = whole.loss_grad_and_filter(ts[k], ps[k], post, 0.1)
        det_w = whole.loss_detail.clone()
        grad_s = None
        for phases in (1, 2, 8, 4):
            tot_s, grad_s, _, _ = split.loss_grad_and_filter(ts[k], ps[k], post, 0.1, phases=phases)
        assert torch.equal(grad_w, grad_s) and torch.equal(det_w, split.loss_detail)
        assert whole.current_num == split.current_num == n
        assert (float(det_w[5].abs().sum()) > 0) == (k == 0)
