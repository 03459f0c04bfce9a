"""The evaluation-while-training step (loss fwd+bwd + filter + NMS of the same head output) as a three-stream pipeline.

One step is seven kernels: assignment, fused sweep (``yv3_loss_filter_fused``), fold, gather, sort, NMS x2.  Only the
sweep is HBM-bound; the other six are latency-bound and small.  Issued one after the other on one stream a step costs
the sum; here the sweeps of consecutive steps run back to back on the caller's stream and the small kernels run under
them on two side streams:

    main   sweep(k) ............ sweep(k+1) ............ sweep(k+2)
    aux    fold(k-1) assign(k+depth-1)   fold(k) assign(k+depth)
    post   gather(k-1) sort NMS(k-1)     gather(k) sort NMS(k)          (high priority)

``depth`` (loss workspace, post-process workspace) pairs are used round-robin.  The edges that keep it correct:

  * sweep(k)   waits for assign(k)                       (object bits, label slots, owner terms of step k)
  * sweep(k)   waits for gather(k - depth)               (it rewrites the staging rows that gather reads)
  * fold(k), gather(k) wait for sweep(k)
  * assign(k + depth) is queued behind fold(k) on aux    (it rewrites the loss workspace step k used)
  * gather/sort/NMS of consecutive steps share cand / counts / keep only when the caller passes single buffers; they
    are ordered by the post stream itself

``tests/test_gpu_pipeline.py`` runs K distinct batches through it, eagerly and as a CUDA-graph replay, and compares every
step's loss row, gradient, candidates, counts and keep flags bit for bit with the serial single-stream calls.

The rectified-term switch of yolov3_loss.py:125-134 is ONE counter for the whole pipeline: the flag of step k is fixed
when the step is issued (``current_num`` advances by the global batch per active step), whichever workspace pair runs it.
"""
import torch

from .yolov3_loss import YOLOv3Loss
from .yolov3_post_process import HeadPostProcessor


def _nth(x, k):
    return x[k] if isinstance(x, (list, tuple)) else x


class FusedPipeline(object):
    def __init__(self, loss_args, score_thresh, nms_thresh, depth=3, rectified_coord_num=0, process_group=None,
                 high_priority=True, reducer=None):
        """``loss_args``: keyword arguments of YOLOv3Loss (without rectified_coord_num).  ``reducer``: a stream-ordered
        in-place sum over the ranks of one 20-float row (``PeerAllReduce``); when given, every step's row is reduced
        INSIDE the step, on the side stream right behind its fold (and inside a captured graph); without it the caller
        reduces the collected rows with ``all_reduce_rows`` (one NCCL call for all of them)."""
        if depth < 1:
            raise ValueError("depth must be >= 1")
        self.score_thresh, self.nms_thresh = score_thresh, nms_thresh
        self.rectified_coord_num = int(rectified_coord_num)
        self.current_num = 0
        self.pairs = []
        for _ in range(depth):
            lo = YOLOv3Loss(rectified_coord_num=rectified_coord_num, process_group=process_group, **loss_args)
            lo.auto_all_reduce = False              # the pipeline's caller reduces the rows it collected (all_reduce_rows)
            po = HeadPostProcessor(lo.head_grid_sizes, lo.class_num, loss_args["anchor_boxes"])
            self.pairs.append((lo, po))
        self.process_group = process_group
        self.reducer = reducer
        self.capacity = self.pairs[0][1].capacity
        prio = -1 if high_priority else 0
        self.aux_stream = torch.cuda.Stream(priority=prio)
        self.post_stream = torch.cuda.Stream(priority=prio)
        self.reduce_stream = torch.cuda.Stream(priority=prio)
        self.last_fused = None

    # ---- bookkeeping ---------------------------------------------------------------------------------------------
    def _rect_flags(self, count, n_global):
        """tf.cond(current_num <= rectified_coord_num) per step, read before that step's update_add (:125, :152)."""
        flags, cur = [], self.current_num
        for _ in range(count):
            act = cur <= self.rectified_coord_num
            flags.append(act)
            if act:
                cur += n_global
        return flags, cur

    def state_dict(self):
        return {"current_num": int(self.current_num)}

    def load_state_dict(self, state):
        self.current_num = int(state["current_num"])

    def check_status(self):
        for lo, po in self.pairs:
            lo.check_status()
            po.check_status()

    # ---- the schedule ----------------------------------------------------------------------------------------------
    def run(self, targets, predicts, grads, rows, cand, counts, keep, count=None, upstream=1.0, serial=False):
        """``count`` steps.  Every argument is either one tensor used by all steps or a list with one entry per step:
        targets (N, M*5), predicts (packed), grads (like predicts), cand (N, capacity, 8), counts (N, 3), keep (N, capacity);
        ``rows`` is a (count, 20) tensor: row k receives step k's [18 detail values, their sum, 0] (this rank's share
        when the batch is sharded: see all_reduce_rows).  ``serial=True`` issues the same calls one after the other on the
        current stream with the first workspace pair (the reference schedule the tests compare with).
        Returns nothing; everything is stream-ordered behind the current stream when the call returns."""
        if count is None:
            count = len(predicts) if isinstance(predicts, (list, tuple)) else rows.shape[0]
        if rows.shape[0] < count or rows.shape[1] != 20:
            raise ValueError("rows must be (count, 20)")
        p0 = _nth(predicts, 0)
        lo0 = self.pairs[0][0]
        world = lo0._world()
        n_global = lo0._n_global(p0.shape[0], world, p0.device)
        flags, cur_after = self._rect_flags(count, n_global)
        depth = len(self.pairs)
        thr = self.score_thresh
        cur = torch.cuda.current_stream(p0.device)

        def part(k, phases, pair=None):
            lo, po = self.pairs[k % depth] if pair is None else pair
            return lo.loss_grad_and_filter(_nth(targets, k), _nth(predicts, k), po, thr, upstream=upstream,
                                           grad_out=_nth(grads, k), out=rows[k],
                                           filter_out=(_nth(cand, k), _nth(counts, k)), phases=phases,
                                           rect_override=flags[k])

        if serial:
            for k in range(count):
                part(k, 15, self.pairs[0])
                if self.reducer is not None:
                    self.reducer(rows[k])
                self.pairs[0][1].nms(_nth(cand, k), _nth(counts, k), self.nms_thresh, out=_nth(keep, k))
            self.last_fused = self.pairs[0][0].last_fused
            self._finish(cur_after)
            return

        aux, post = self.aux_stream, self.post_stream

        def assign(k):
            with torch.cuda.stream(aux):
                part(k, 1)
                e = torch.cuda.Event()
                e.record(aux)
            return e

        aux.wait_stream(cur)
        post.wait_stream(cur)
        if self.reducer is not None:
            self.reduce_stream.wait_stream(cur)
        a_done = [assign(j) if j < count else None for j in range(depth)]
        g_done = [None] * depth
        for k in range(count):
            b = k % depth
            cur.wait_event(a_done[b])
            if g_done[b] is not None:
                cur.wait_event(g_done[b])
            part(k, 2)
            swept = torch.cuda.Event()
            swept.record(cur)
            with torch.cuda.stream(post):
                post.wait_event(swept)
                part(k, 8)
                self.pairs[b][1].nms(_nth(cand, k), _nth(counts, k), self.nms_thresh, out=_nth(keep, k))
                g_done[b] = torch.cuda.Event()
                g_done[b].record(post)
            with torch.cuda.stream(aux):
                aux.wait_event(swept)
                part(k, 4)
                if self.reducer is not None:
                    folded = torch.cuda.Event()
                    folded.record(aux)
            if self.reducer is not None:
                # the step's one exchange: 80 bytes over NVLink, on a stream of its own -- nothing of this rank's later
                # steps waits for the peers (behind the fold on the aux stream it held back assign(k + depth), and with it
                # sweep k + depth, until the slowest rank had folded step k: the ranks ran in lock step)
                with torch.cuda.stream(self.reduce_stream):
                    self.reduce_stream.wait_event(folded)
                    self.reducer(rows[k])
            if k + depth < count:
                a_done[b] = assign(k + depth)
        cur.wait_stream(post)
        cur.wait_stream(aux)
        if self.reducer is not None:
            cur.wait_stream(self.reduce_stream)
        self.last_fused = all(lo.last_fused for lo, _ in self.pairs[:min(depth, count)])
        self._finish(cur_after)

    def _finish(self, cur_after):
        self.current_num = cur_after
        for lo, _ in self.pairs:
            lo.current_num = cur_after

    def all_reduce_rows(self, rows):
        """Sum the loss rows of the steps just issued over the ranks: ONE collective for all of them (80 bytes per step).
        Stream-ordered on the current stream."""
        lo = self.pairs[0][0]
        if self.reducer is None and lo._world() > 1:
            lo._all_reduce(rows)
        return rows

    def capture(self, targets, predicts, grads, rows, cand, counts, keep, count=None, upstream=1.0):
        """The same ``run`` captured once in a CUDA graph (the buffers are baked in; refill them and replay).
        The rectified switch is a kernel argument, so a captured graph holds the flags of the steps it was captured
        for: capture when the correction window is closed or far from closing, or re-capture when it closes.
        Returns the torch.cuda.CUDAGraph."""
        state = self.current_num
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):                  # warm-up outside the capture (workspaces, attributes)
            self.run(targets, predicts, grads, rows, cand, counts, keep, count=count, upstream=upstream)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self._finish(state)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            self.run(targets, predicts, grads, rows, cand, counts, keep, count=count, upstream=upstream)
        self._finish(state)
        return graph
