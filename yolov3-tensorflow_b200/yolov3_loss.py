"""Drop-in for yolov3/yolov3_loss.py::YOLOv3Loss backed by yv3_loss_fwd_bwd (CUDA, sm_100a).

The reference builds the loss as ~150 small TF ops per image inside tf.map_fn and lets autodiff derive
the backward; here one assignment kernel plus one fused sweep produce the loss detail AND
d loss / d predicts in a single pass, so ``loss()`` hands torch.autograd a ready gradient.
"""
import ctypes

import numpy as np
import torch

from . import _lib

_DETAIL_NAMES = ("coord_loss_xy", "coord_loss_wh", "noobj_iou_loss", "obj_iou_loss", "class_loss",
                 "rectified_coord_loss")


class _FusedLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, predicts, owner, targets):
        total, grad = owner._run(targets, predicts, want_grad=True)
        ctx.save_for_backward(grad)
        return total

    @staticmethod
    def backward(ctx, grad_output):
        (grad,) = ctx.saved_tensors
        # upstream is 1 under Keras (the loss IS the objective, trainer.py:84) but it arrives as a device scalar: reading
        # it would stall the host, so the saved gradient is scaled in place (one pass, no second tensor) -- it is this
        # function's own buffer, nobody else holds it.  loss_and_grad(upstream=...) scales inside the kernel instead.
        return grad.mul_(grad_output), None, None


class YOLOv3Loss(object):
    """Same constructor arguments and ``loss(targets, predicts)`` contract as the reference
    (yolov3_loss.py:16-18, :81).  Side contract kept: the six ``loss_detail`` vectors (shape (3,),
    heads /8,/16,/32) are refreshed by every call, and ``current_num`` counts the images seen while the
    train-from-scratch correction term is active (:69, :125-134, :152).

    Multi-GPU: one process per GPU, each holding a contiguous shard of the batch.  When
    torch.distributed is initialised (or ``process_group`` is given) the batch mean runs over the
    GLOBAL batch: gradients are scaled by 1/N_global locally and the 20 loss floats are all-reduced
    (NCCL); no other data crosses NVLink."""

    def __init__(self, head_grid_sizes, class_num, anchor_boxes, iou_thresh, loss_weights,
                 rectified_coord_num=0, rectified_loss_weight=None,
                 is_focal_loss=False, focal_alpha=0.25, focal_gamma=2.0, is_tiou_recall=False,
                 process_group=None):
        self.head_grid_sizes = [(int(h), int(w)) for (h, w) in head_grid_sizes]
        self.class_num = int(class_num)
        self.box_len = 5 + self.class_num
        self.iou_thresh = iou_thresh
        self.is_focal_loss = is_focal_loss
        self.focal_alpha = focal_alpha
        self.focal_gamma = focal_gamma
        self.is_tiou_recall = is_tiou_recall
        self.rectified_coord_num = int(rectified_coord_num)
        self._box_num = [len(a) for a in anchor_boxes]
        self._cfg = _lib.make_config(self.head_grid_sizes, self.class_num, anchor_boxes, iou_thresh,
                                     loss_weights, rectified_loss_weight, is_focal_loss, focal_alpha,
                                     focal_gamma, is_tiou_recall)
        self._lib = _lib.lib()
        self.records_per_image = int(self._lib.yv3_records_per_image(ctypes.byref(self._cfg)))
        self.packed_floats_per_image = int(self._lib.yv3_packed_floats_per_image(ctypes.byref(self._cfg)))
        self.process_group = process_group
        # False: the caller sums the 20 loss floats over the ranks itself with all_reduce_detail() (e.g. after replaying
        # a CUDA graph that holds this rank's launches)
        self.auto_all_reduce = True
        # Equal shards are assumed (n_global = n * world) and verified ONCE, on the first sharded call, by an all-reduce
        # of the local batch size; a caller with ragged shards (a last, shorter batch) passes ``global_batch``.
        self.global_batch = None
        self._shards_checked = False
        self._latched_rect = None           # the tf.cond of a step, latched by the step's first phase call
        self._last_out = None
        # host-visible state of the tf.cond at yolov3_loss.py:125 (a bare backend variable there, :69)
        self.current_num = 0
        self._ws = {}
        self._last_ws = None
        self.loss_detail = None
        for name in _DETAIL_NAMES:
            setattr(self, name, None)

    # ---- distributed plumbing --------------------------------------------------------------------
    def _world(self):
        import torch.distributed as dist
        if self.process_group is not None:
            return dist.get_world_size(self.process_group)
        if dist.is_available() and dist.is_initialized():
            return dist.get_world_size()
        return 1

    def _all_reduce(self, t):
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.process_group)

    def _n_global(self, n, world, device):
        """The N of the batch mean (yolov3_loss.py:112) when the batch is sharded over ranks."""
        if self.global_batch is not None:
            return int(self.global_batch)
        if world > 1 and not self._shards_checked:
            import torch.distributed as dist
            v = torch.tensor([n, -n], dtype=torch.int64, device=device)
            dist.all_reduce(v, op=dist.ReduceOp.MAX, group=self.process_group)
            if int(v[0]) != -int(v[1]):
                raise ValueError("ranks hold shards of different sizes (%d..%d images): set loss.global_batch to the "
                                 "global batch size; n * world would give a wrong batch mean" % (-int(v[1]), int(v[0])))
            self._shards_checked = True
        return n * world

    # ---- the fused call --------------------------------------------------------------------------
    def _workspace(self, n, m, device):
        key = (n, m, device)
        ws = self._ws.get(key)
        if ws is None:
            size = ctypes.c_size_t(0)
            _lib.check(self._lib.yv3_query_workspace(ctypes.byref(self._cfg), n, m, ctypes.byref(size)),
                       "yv3_query_workspace")
            ws = torch.empty(size.value, dtype=torch.uint8, device=device)
            self._ws[key] = ws
        return ws

    def _prepare(self, targets, predicts):
        p = _lib.to_device_f32(predicts)
        t = _lib.to_device_f32(targets, p.device)
        if p.dim() != 4 or p[0].numel() != self.packed_floats_per_image:
            raise ValueError("predicts must be (N, H/32, W/32, %d floats per image in total), got %s" %
                             (self.packed_floats_per_image, tuple(p.shape)))
        if t.dim() != 2 or t.shape[0] != p.shape[0] or t.shape[1] % 5 != 0 or t.shape[1] == 0:
            raise ValueError("targets must be (N, M*5) with the same N as predicts")
        return t, p

    def _run(self, targets, predicts, want_grad, upstream=1.0, events=None, grad_out=None, sweep_after=None,
             sweep_done=None, out=None, fuse=None, rect_override=None):
        t, p = self._prepare(targets, predicts)
        with torch.cuda.device(p.device):           # launches go to the tensors' device, whatever the current one is
            return self._run_on(t, p, want_grad, upstream, events, grad_out, sweep_after, sweep_done, out, fuse,
                                rect_override)

    def _run_on(self, t, p, want_grad, upstream, events, grad_out, sweep_after, sweep_done, out, fuse, rect_override):
        n, m = p.shape[0], t.shape[1] // 5
        world = self._world()
        n_global = self._n_global(n, world, p.device)
        # :125 -- the predicate is read once per step, before the add.  A scheduler that issues the phases of a step
        # as separate calls gets the value latched by the step's first call (the counter moves in the call that holds
        # the fold); a pipeline that interleaves steps of several instances passes it explicitly (rect_override).
        if rect_override is not None:
            rect_active = bool(rect_override)
        elif self._latched_rect is not None:
            rect_active = self._latched_rect
        else:
            rect_active = self.current_num <= self.rectified_coord_num
        ws = self._workspace(n, m, p.device)
        if out is None:
            out = torch.empty(20, dtype=torch.float32, device=p.device)
        elif out.shape != (20,) or out.dtype != torch.float32 or not out.is_contiguous() or out.device != p.device:
            raise ValueError("out must be a contiguous fp32 tensor of 20 elements on the device of predicts")
        grad = None
        if want_grad:
            grad = torch.empty_like(p) if grad_out is None else grad_out
            if grad.shape != p.shape or grad.dtype != torch.float32 or not grad.is_contiguous() or grad.device != p.device:
                raise ValueError("grad_out must be a contiguous fp32 tensor shaped like predicts on the same device")
        args = (ctypes.byref(self._cfg), _lib.ptr(p), _lib.ptr(t), n, m, n_global, 1 if rect_active else 0,
                float(upstream), _lib.ptr(out), _lib.ptr(grad), _lib.ptr(ws), ws.numel())
        stream = _lib.current_stream_ptr(p.device)
        if fuse is not None:
            # the loss and the post-process filter of the same predicts on one read of the tensor (fused_sweep.cu)
            post, score_thresh, filter_out, before_tail, phases = fuse
            if post.packed_floats_per_image != self.packed_floats_per_image:
                raise ValueError("the post-processor was built for another head configuration")
            pws = post._workspace(n, p.device)
            if filter_out is None:
                cand = torch.empty((n, post.capacity, 8), dtype=torch.float32, device=p.device)
                counts = torch.empty((n, 3), dtype=torch.int32, device=p.device)
            else:
                cand, counts = filter_out
            fused = ctypes.c_int32(0)
            fargs = args + (float(np.float32(score_thresh)), _lib.ptr(cand), post.capacity, _lib.ptr(counts),
                            _lib.ptr(pws), pws.numel(), ctypes.byref(fused))
            if phases != 15:
                _lib.check(self._lib.yv3_loss_filter_fused(*fargs, phases, stream), "yv3_loss_filter_fused[%d]" % phases)
            elif events is None and before_tail is None:
                _lib.check(self._lib.yv3_loss_filter_fused(*fargs, 15, stream), "yv3_loss_filter_fused")
            else:
                # (start, stop) events around the sweep alone; `before_tail`: an event (of another stream) the fold +
                # gather tail waits for, e.g. the previous step's NMS still reading cand / counts
                cur = torch.cuda.current_stream(p.device)
                _lib.check(self._lib.yv3_loss_filter_fused(*fargs, 1, stream), "yv3_loss_filter_fused[assign]")
                if events is not None:
                    events[0].record()
                _lib.check(self._lib.yv3_loss_filter_fused(*fargs, 2, stream), "yv3_loss_filter_fused[sweep]")
                if events is not None:
                    events[1].record()
                if before_tail is not None:
                    cur.wait_event(before_tail)
                _lib.check(self._lib.yv3_loss_filter_fused(*fargs, 12, stream), "yv3_loss_filter_fused[tail]")
            post._last_ws = pws
            self.last_fused = bool(fused.value)
            self._last_filter = (cand, counts)
            if not phases & 4:
                # a scheduler issuing the phases one by one: the step's bookkeeping belongs to the call with the fold
                if rect_override is None:
                    self._latched_rect = rect_active
                return None, grad
        elif sweep_after is not None or sweep_done is not None:
            # a scheduler's hooks around the HBM-bound sweep: it starts after `sweep_after` (an event of another
            # stream) and `sweep_done` is recorded right behind it; the small assignment and fold kernels are free
            cur = torch.cuda.current_stream(p.device)
            _lib.check(self._lib.yv3_loss_fwd_bwd(*args, 1, stream), "yv3_loss_fwd_bwd[assign]")
            if sweep_after is not None:
                cur.wait_event(sweep_after)
            _lib.check(self._lib.yv3_loss_fwd_bwd(*args, 2, stream), "yv3_loss_fwd_bwd[dense]")
            if sweep_done is not None:
                sweep_done.record(cur)
            _lib.check(self._lib.yv3_loss_fwd_bwd(*args, 4, stream), "yv3_loss_fwd_bwd[fold]")
        elif events is None:
            _lib.check(self._lib.yv3_loss_fwd_bwd(*args, 7, stream), "yv3_loss_fwd_bwd")
        else:
            # (start, stop) torch.cuda.Event pair bracketing the dense sweep alone (bench.py's roofline)
            _lib.check(self._lib.yv3_loss_fwd_bwd(*args, 1, stream), "yv3_loss_fwd_bwd[assign]")
            events[0].record()
            _lib.check(self._lib.yv3_loss_fwd_bwd(*args, 2, stream), "yv3_loss_fwd_bwd[dense]")
            events[1].record()
            _lib.check(self._lib.yv3_loss_fwd_bwd(*args, 4, stream), "yv3_loss_fwd_bwd[fold]")
        self._latched_rect = None
        if rect_active and rect_override is None:
            self.current_num += n_global                                       # update_add (:152)
        if world > 1 and self.auto_all_reduce:
            self._all_reduce(out)
        self._last_out = out
        self._last_ws = ws
        detail = out[:18].view(6, 3)
        self.loss_detail = detail
        for i, name in enumerate(_DETAIL_NAMES):
            setattr(self, name, detail[i])
        return out[18], grad

    # ---- public surface --------------------------------------------------------------------------
    def loss(self, targets, predicts):
        """Scalar total loss (a 0-dim CUDA tensor).  Differentiable with respect to ``predicts`` when it
        requires grad: the gradient was produced by the same fused pass."""
        if isinstance(predicts, torch.Tensor) and predicts.requires_grad:
            return _FusedLoss.apply(predicts, self, targets)
        total, _ = self._run(targets, predicts, want_grad=False)
        return total

    __call__ = loss

    def loss_and_grad(self, targets, predicts, upstream=1.0, events=None, grad_out=None, sweep_after=None,
                      sweep_done=None, out=None):
        """(total loss, d(upstream * loss)/d predicts in the packed layout) from one fused pass.
        ``grad_out`` reuses a caller-owned gradient buffer; ``events`` = (start, stop) CUDA events recorded
        around the dense sweep alone; ``out`` = a caller-owned 20-float slot for [18 detail values, their sum, 0] (e.g.
        row k of a (steps, 20) buffer that is all-reduced once for several steps)."""
        return self._run(targets, predicts, want_grad=True, upstream=upstream, events=events, grad_out=grad_out,
                         sweep_after=sweep_after, sweep_done=sweep_done, out=out)

    def loss_grad_and_filter(self, targets, predicts, post, score_thresh, upstream=1.0, grad_out=None, out=None,
                             filter_out=None, events=None, before_tail=None, phases=15, rect_override=None):
        """Evaluation while training: ``loss_and_grad`` and ``post.filter(predicts, score_thresh)`` (a
        ``HeadPostProcessor`` of the same head configuration) in one call that reads the head tensor once
        (``yv3_loss_filter_fused``).  Returns (total, grad, cand, counts); gradient, candidates and counts are
        bit-identical to the two separate calls.  ``self.last_fused`` tells whether the one-sweep kernel ran or the call
        fell back to the two sweeps (unaligned tensor, more than 262144 tiles).  ``post.nms(cand, counts, thr)``
        follows as usual.  ``events`` = (start, stop) CUDA events recorded around the sweep alone; ``before_tail`` = an
        event the fold + gather tail waits for (a scheduler's hook: the previous step's NMS on another stream).
        ``phases`` (1 assignment, 2 sweep, 4 fold, 8 gather; 15 = all): a scheduler may issue the parts as separate
        calls with the same arguments on streams of its choice (bench.py runs the sweeps back to back on one stream and
        the small kernels on two others, two workspaces alternating); total is None unless the call includes the fold,
        which is also the call that does the step's bookkeeping (rectified counter, all-reduce, detail attributes)."""
        total, grad = self._run(targets, predicts, want_grad=True, upstream=upstream, grad_out=grad_out, out=out,
                                events=events, fuse=(post, score_thresh, filter_out, before_tail, int(phases)),
                                rect_override=rect_override)
        cand, counts = self._last_filter
        return total, grad, cand, counts

    def loss_and_grad_heads(self, targets, head_predicts, upstream=1.0, want_grad=True, grads_out=None):
        """The same loss on the three head tensors BEFORE yolov3_detector.py:80-85 packs them: ``head_predicts`` =
        [/8, /16, /32] conv outputs, each (N, H_h, W_h, B_h * (5 + C)) (or (..., B_h, 5 + C)) fp32 contiguous.
        Returns (total loss, [gradient per head tensor, same shapes] or None).  Saves the Reshape + concatenate copy
        on the way in and the split of the gradient on the way out."""
        if len(head_predicts) != 3:
            raise ValueError("head_predicts must hold the /8, /16 and /32 head tensors")
        heads = [_lib.to_device_f32(p) for p in head_predicts]
        t = _lib.to_device_f32(targets, heads[0].device)
        n = heads[0].shape[0]
        for h, p in enumerate(heads):
            gh, gw = self.head_grid_sizes[h]
            if p.shape[0] != n or tuple(p.shape[1:3]) != (gh, gw) or p[0].numel() != gh * gw * self._box_num[h] * self.box_len:
                raise ValueError("head %d must be (N, %d, %d, %d), got %s" % (
                    h, gh, gw, self._box_num[h] * self.box_len, tuple(p.shape)))
        if t.dim() != 2 or t.shape[0] != n or t.shape[1] % 5 != 0 or t.shape[1] == 0:
            raise ValueError("targets must be (N, M*5) with the same N as the head tensors")
        m = t.shape[1] // 5
        world = self._world()
        n_global = self._n_global(n, world, heads[0].device)
        rect_active = self.current_num <= self.rectified_coord_num            # :125 (read before the add)
        ws = self._workspace(n, m, heads[0].device)
        out = torch.empty(20, dtype=torch.float32, device=heads[0].device)
        grads = None
        if want_grad:
            grads = [torch.empty_like(p) for p in heads] if grads_out is None else list(grads_out)
            for g, p in zip(grads, heads):
                if g.shape != p.shape or g.dtype != torch.float32 or not g.is_contiguous() or g.device != p.device:
                    raise ValueError("grads_out must be contiguous fp32 tensors shaped like the head tensors")
        pp = (ctypes.c_void_p * 3)(*[_lib.ptr(p) for p in heads])
        gp = (ctypes.c_void_p * 3)(*[_lib.ptr(g) for g in grads]) if grads is not None else None
        with torch.cuda.device(heads[0].device):
            rc = self._lib.yv3_loss_fwd_bwd_heads(ctypes.byref(self._cfg), pp, _lib.ptr(t), n, m, n_global,
                                                  1 if rect_active else 0, float(upstream), _lib.ptr(out), gp, _lib.ptr(ws),
                                                  ws.numel(), 7, _lib.current_stream_ptr(heads[0].device))
        _lib.check(rc, "yv3_loss_fwd_bwd_heads")
        if rect_active:
            self.current_num += n_global                                       # update_add (:152)
        if world > 1 and self.auto_all_reduce:
            self._all_reduce(out)
        self._last_out = out
        self._last_ws = ws
        detail = out[:18].view(6, 3)
        self.loss_detail = detail
        for i, name in enumerate(_DETAIL_NAMES):
            setattr(self, name, detail[i])
        return out[18], grads

    def loss_grad_and_filter_heads(self, targets, head_predicts, post, score_thresh, upstream=1.0, grads_out=None,
                                   filter_out=None, phases=15):
        """``loss_grad_and_filter`` on the three head tensors BEFORE yolov3_detector.py:80-85 packs them
        (``yv3_loss_filter_fused_heads``): every head tensor is read once, nothing is concatenated.  Returns
        (total, [gradient per head tensor], cand, counts); bit-identical to ``loss_and_grad_heads`` + ``post.filter_heads``.
        ``self.last_fused`` tells whether the one-sweep kernel ran."""
        if len(head_predicts) != 3:
            raise ValueError("head_predicts must hold the /8, /16 and /32 head tensors")
        heads = [_lib.to_device_f32(p) for p in head_predicts]
        dev = heads[0].device
        t = _lib.to_device_f32(targets, dev)
        n = heads[0].shape[0]
        for h, p in enumerate(heads):
            gh, gw = self.head_grid_sizes[h]
            if p.shape[0] != n or tuple(p.shape[1:3]) != (gh, gw) or p[0].numel() != gh * gw * self._box_num[h] * self.box_len:
                raise ValueError("head %d must be (N, %d, %d, %d), got %s" % (
                    h, gh, gw, self._box_num[h] * self.box_len, tuple(p.shape)))
        if t.dim() != 2 or t.shape[0] != n or t.shape[1] % 5 != 0 or t.shape[1] == 0:
            raise ValueError("targets must be (N, M*5) with the same N as the head tensors")
        if post.packed_floats_per_image != self.packed_floats_per_image:
            raise ValueError("the post-processor was built for another head configuration")
        m = t.shape[1] // 5
        world = self._world()
        n_global = self._n_global(n, world, dev)
        rect_active = self._latched_rect if self._latched_rect is not None else self.current_num <= self.rectified_coord_num
        ws = self._workspace(n, m, dev)
        pws = post._workspace(n, dev)
        out = torch.empty(20, dtype=torch.float32, device=dev)
        grads = [torch.empty_like(p) for p in heads] if grads_out is None else list(grads_out)
        for g, p in zip(grads, heads):
            if g.shape != p.shape or g.dtype != torch.float32 or not g.is_contiguous() or g.device != p.device:
                raise ValueError("grads_out must be contiguous fp32 tensors shaped like the head tensors")
        if filter_out is None:
            cand = torch.empty((n, post.capacity, 8), dtype=torch.float32, device=dev)
            counts = torch.empty((n, 3), dtype=torch.int32, device=dev)
        else:
            cand, counts = filter_out
        pp = (ctypes.c_void_p * 3)(*[_lib.ptr(p) for p in heads])
        gp = (ctypes.c_void_p * 3)(*[_lib.ptr(g) for g in grads])
        fused = ctypes.c_int32(0)
        with torch.cuda.device(dev):
            rc = self._lib.yv3_loss_filter_fused_heads(
                ctypes.byref(self._cfg), pp, _lib.ptr(t), n, m, n_global, 1 if rect_active else 0, float(upstream), _lib.ptr(out),
                gp, _lib.ptr(ws), ws.numel(), float(np.float32(score_thresh)), _lib.ptr(cand), post.capacity, _lib.ptr(counts),
                _lib.ptr(pws), pws.numel(), ctypes.byref(fused), int(phases), _lib.current_stream_ptr(dev))
        _lib.check(rc, "yv3_loss_filter_fused_heads")
        post._last_ws = pws
        self.last_fused = bool(fused.value)
        self._last_filter = (cand, counts)
        self._last_ws = ws
        if not phases & 4:
            self._latched_rect = rect_active
            return None, grads, cand, counts
        self._latched_rect = None
        if rect_active:
            self.current_num += n_global                                       # update_add (:152)
        if world > 1 and self.auto_all_reduce:
            self._all_reduce(out)
        self._last_out = out
        detail = out[:18].view(6, 3)
        self.loss_detail = detail
        for i, name in enumerate(_DETAIL_NAMES):
            setattr(self, name, detail[i])
        return out[18], grads, cand, counts

    def all_reduce_detail(self):
        """Sum the last call's 20 loss floats over the ranks (only needed with ``auto_all_reduce = False``)."""
        if self._last_out is not None and self._world() > 1:
            self._all_reduce(self._last_out)

    def assign(self, targets, predicts):
        """The integer decisions of the loss: (table (N,M,3,4) int32 [row, col, best_anchor, owned] per
        label slot and head, -1/-1/-1/0 for padding; iou (N,M,3) fp32 best centre-cell IoU)."""
        t, p = self._prepare(targets, predicts)
        n, m = p.shape[0], t.shape[1] // 5
        ws = self._workspace(n, m, p.device)
        table = torch.empty((n, m, 3, 4), dtype=torch.int32, device=p.device)
        iou = torch.empty((n, m, 3), dtype=torch.float32, device=p.device)
        with torch.cuda.device(p.device):
            rc = self._lib.yv3_assign(ctypes.byref(self._cfg), _lib.ptr(p), _lib.ptr(t), n, m, _lib.ptr(table),
                                      _lib.ptr(iou), _lib.ptr(ws), ws.numel(), _lib.current_stream_ptr(p.device))
        _lib.check(rc, "yv3_assign")
        self._last_ws = ws
        return table, iou

    def check_status(self):
        """Synchronises and raises what TensorFlow would have raised for the last call: IndexError for a
        label centre outside the grid (gather_nd, yolov3_loss.py:271); FloatingPointError for inf/NaN
        in the loss (which TF silently propagates)."""
        if self._last_ws is None:
            return
        rc = self._lib.yv3_read_status(_lib.ptr(self._last_ws),
                                       _lib.current_stream_ptr(self._last_ws.device))
        if rc == _lib.YV3_ERR_GT_RANGE:
            raise IndexError(_lib.strerror(rc))
        if rc == _lib.YV3_ERR_NONFINITE:
            raise FloatingPointError(_lib.strerror(rc))
        _lib.check(rc, "yv3_read_status")

    def state_dict(self):
        """``current_num`` is not checkpointed by the reference (it is not a model weight); exposing it
        lets a trainer resume the correction phase where it stopped."""
        return {"current_num": int(self.current_num)}

    def load_state_dict(self, state):
        self.current_num = int(state["current_num"])
