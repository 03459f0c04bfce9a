"""GPU: stand-in for compute-sanitizer, which is closed on this pool (profiles/r02_sanitizer.txt): guard bands around
every output and workspace (out-of-bounds writes), poisoned initial contents (reads of uninitialised memory) and
repeated runs (races) for the loss, fused, un-packed, filter, NMS, collect and mAP entry points."""
import numpy as np
import pytest
import torch

from helpers import make_predicts, make_spec, make_targets, product_loss_args

pytestmark = pytest.mark.gpu
GUARD = 4096          # bytes on each side


@pytest.fixture(scope="module")
def yv3():
    import yolov3_tensorflow_b200 as pkg
    from yolov3_tensorflow_b200 import _lib
    _lib.lib()
    return pkg


class Guarded(object):
    """tensors carved out of one byte buffer with canary bands between them"""

    def __init__(self, poison):
        self.items, self.poison = [], poison

    def make(self, shape, dtype):
        nbytes = int(np.prod(shape)) * torch.empty((), dtype=dtype).element_size()
        pad = (-nbytes) % 256
        buf = torch.full((GUARD + nbytes + pad + GUARD,), 0xA5, dtype=torch.uint8, device="cuda")
        body = buf[GUARD:GUARD + nbytes]
        body.fill_(self.poison)
        self.items.append((buf, nbytes))
        return body.view(dtype).view(shape)

    def check(self):
        torch.cuda.synchronize()
        for buf, nbytes in self.items:
            assert bool((buf[:GUARD] == 0xA5).all()), "a kernel wrote BEFORE one of its buffers"
            assert bool((buf[GUARD + nbytes + ((-nbytes) % 256):] == 0xA5).all()), "a kernel wrote PAST one of its buffers"


def _run_all(yv3, spec, p, t, thr, poison):
    """every call once, all outputs and workspaces guarded; returns the results to compare across runs"""
    n, m = p.shape[0], t.shape[1] // 5
    gd = Guarded(poison)
    post = yv3.HeadPostProcessor(spec.head_grid_sizes, spec.class_num, spec.anchor_boxes)
    loss = yv3.YOLOv3Loss(rectified_coord_num=10 ** 9, **product_loss_args(spec))
    # workspaces: ask the wrappers for their sizes, then swap guarded ones in
    ws = loss._workspace(n, m, p.device)
    loss._ws[(n, m, p.device)] = gd.make((ws.numel(),), torch.uint8)
    pws = post._workspace(n, p.device)
    post._ws[(n, p.device)] = gd.make((pws.numel(),), torch.uint8)
    cap = post.capacity
    out = {}
    grad = gd.make(tuple(p.shape), torch.float32)
    row = gd.make((20,), torch.float32)
    loss.loss_and_grad(t, p, grad_out=grad, out=row)
    out["dense"] = (grad.clone(), row.clone())
    grad2 = gd.make(tuple(p.shape), torch.float32)
    cand = gd.make((n, cap, 8), torch.float32)
    counts = gd.make((n, 3), torch.int32)
    row2 = gd.make((20,), torch.float32)
    loss.loss_grad_and_filter(t, p, post, thr, grad_out=grad2, out=row2, filter_out=(cand, counts))
    keep = gd.make((n, cap), torch.uint8)
    post.nms(cand, counts, 0.45, out=keep)
    tot = counts.sum(dim=1)
    out["fused"] = (grad2.clone(), row2.clone(), counts.clone(),
                    [cand[i, :int(tot[i])].clone() for i in range(n)], [keep[i, :int(tot[i])].clone() for i in range(n)])
    heads = [p[..., spec.chan_base[h]:spec.chan_base[h] + spec.head_channels[h]].reshape(
        n, spec.head_grid_sizes[h][0], spec.head_grid_sizes[h][1], -1).contiguous() for h in range(3)]
    hg = [gd.make(tuple(h.shape), torch.float32) for h in heads]
    cand3 = gd.make((n, cap, 8), torch.float32)
    counts3 = gd.make((n, 3), torch.int32)
    loss.loss_grad_and_filter_heads(t, heads, post, thr, grads_out=hg, filter_out=(cand3, counts3))
    out["heads"] = ([g.clone() for g in hg], counts3.clone())
    H, W = spec.head_grid_sizes[0][0] * 8, spec.head_grid_sizes[0][1] * 8
    rows, kc = post.collect(dict(cand=cand, counts=counts, keep=keep), (W, H))
    acc = yv3.MeanAPAccumulator(spec.class_num, (W, H), max_detections=1 << 14)
    acc._state = gd.make((acc._state.numel(),), torch.uint8)
    acc.reset()
    acc.update_rows(rows, kc, t)
    ap, mean, cnt = acc.compute()
    out["map"] = (ap.clone(), mean)
    loss.check_status()
    post.check_status()
    gd.check()
    return out


def _same(a, b):
    if isinstance(a, torch.Tensor):
        return torch.equal(a, b) or (a.is_floating_point() and torch.equal(torch.nan_to_num(a), torch.nan_to_num(b)))
    if isinstance(a, (list, tuple)):
        return len(a) == len(b) and all(_same(x, y) for x, y in zip(a, b))
    return a == b


@pytest.mark.parametrize("name,n,thr", [("coco_64x96_c4", 3, 0.05), ("shipped_384x480_c0", 2, 0.5), ("coco_416_c80", 2, 0.05),
                                        ("single_160_c1_focal", 5, 0.3)])
def test_guard_bands_poison_and_repeats(yv3, name, n, thr):
    spec = make_spec(name)
    yv3.backend.set_epsilon(spec.eps)
    p = torch.from_numpy(make_predicts(spec, n, seed=31)).cuda()
    t = torch.from_numpy(make_targets(spec, n, 7, seed=32, min_valid=0, pad_in_middle=True)).cuda()
    first = _run_all(yv3, spec, p, t, thr, poison=0xFF)          # buffers start as NaN / -1
    again = _run_all(yv3, spec, p, t, thr, poison=0xFF)
    zeros = _run_all(yv3, spec, p, t, thr, poison=0x00)          # ... and as zeros: nothing may depend on the start
    for key in first:
        assert _same(first[key], again[key]), "%s differs between two runs on the same inputs" % key
        assert _same(first[key], zeros[key]), "%s depends on the initial contents of an output or workspace" % key
