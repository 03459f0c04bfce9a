"""Benchmark of the detection-head hot path (BASELINE.json): fused head loss fwd+bwd and decode+filter+NMS.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A step = one pass of the hot path over one synthetic batch: assignment + fused loss forward/backward
(gradient written) followed by fused decode + score threshold + class-aware NMS of the same head tensors.
One JSON line on stdout (rank 0).  See DESIGN.md "Measurement" for every field.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOADS = {
    # BASELINE.json configs[1]: the configuration the metric is quoted on at N=1
    "cfg2_416_c80_b64": dict(size=(416, 416), class_num=80, batch=64, m=16, anchors="coco", iou_thresh=0.8,
                             focal=False, rect=-1, score_thresh=0.1, nms_thresh=0.45,
                             desc="416x416 80-class head (10647 boxes/img), batch 64 per GPU, loss fwd+bwd + "
                                  "decode+filter+NMS"),
    # configs[2] per-GPU shard at 8 GPUs
    "cfg3_608_c80_b64": dict(size=(608, 608), class_num=80, batch=64, m=16, anchors="coco", iou_thresh=0.8,
                             focal=False, rect=-1, score_thresh=0.1, nms_thresh=0.45,
                             desc="608x608 80-class head (22743 boxes/img), batch 64 per GPU"),
    # configs[3]: post-process stress
    "cfg4_416_c80_b256_dense": dict(size=(416, 416), class_num=80, batch=256, m=16, anchors="coco", iou_thresh=0.8,
                                    focal=False, rect=-1, score_thresh=0.001, nms_thresh=0.45,
                                    desc="416x416 batch 256, score thresh 0.001, per-class NMS IoU 0.45 (dense)"),
    # configs[4]: single class, focal + rectified
    "cfg5_512_c0_b16": dict(size=(512, 512), class_num=0, batch=16, m=16, anchors="shipped3", iou_thresh=0.8,
                            focal=True, rect=1 << 30, score_thresh=0.5, nms_thresh=0.4,
                            desc="512x512 single-class head (class_num 0), focal + prior-box correction on, "
                                 "batch 16 per GPU"),
}


def product_spec(w):
    """The workload's switches for the product arm: nothing under oracle/ or tests/ is imported on this path."""
    from yolov3_tensorflow_b200 import synth as ps
    anchors = ps.coco_anchors() if w["anchors"] == "coco" else ps.three_anchor_shipped()
    return ps.ProductSpec(w["size"][0], w["size"][1], w["class_num"], anchors, iou_thresh=w["iou_thresh"],
                          is_focal_loss=w["focal"], focal_alpha=1.0, focal_gamma=2.0, eps=1e-8)


def build_spec(w):
    """The same workload as the oracle's HeadSpec: CPU legs (cpu_baseline, --impl reference) and tests only."""
    from helpers import three_anchor_shipped
    from oracle.head_spec import SHIPPED_LOSS_WEIGHTS, HeadSpec, coco_anchors
    anchors = coco_anchors() if w["anchors"] == "coco" else three_anchor_shipped()
    return HeadSpec.from_input(w["size"][0], w["size"][1], w["class_num"], anchors, iou_thresh=w["iou_thresh"],
                               loss_weights=SHIPPED_LOSS_WEIGHTS, rectified_loss_weight=(1.0, 1.0, 1.0),
                               is_focal_loss=w["focal"], focal_alpha=1.0, focal_gamma=2.0, eps=1e-8)


def synth(spec, n, m, rank):
    """logits iid N(0,1) clipped to [-8,8], seed 1234+rank; labels V~U{1..M}, seed 4321+rank."""
    from yolov3_tensorflow_b200 import synth as ps
    h32, w32 = spec.head_grid_sizes[2]
    return (ps.make_predicts(n, h32 * 32, w32 * 32, spec.box_num, spec.class_num, seed=1234 + rank),
            ps.make_targets(n, spec.class_num, m, seed=4321 + rank, min_valid=1))


class ClockSampler(threading.Thread):
    """nvidia-smi's clocks line through NVML, polled while the benchmark runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag, self.mark = index, [], False, None
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_sm = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.max_sm = None

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self.stop_flag:
            try:
                sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                reasons = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                util = nv.nvmlDeviceGetUtilizationRates(self.h).gpu
                self.samples.append((time.time(), sm, reasons, util, self.mark))
            except Exception:
                pass
            time.sleep(0.005)

    def summary(self):
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_sm, "reasons": ["unavailable"]}
        nv = self.nv
        timed = [s for s in self.samples if s[4] == "timed"] or [s for s in self.samples if s[4] is not None] \
            or self.samples
        names = {
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
            "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
            "hw_power_brake": getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80),
        }
        seen = sorted({k for s in timed for k, bit in names.items() if s[2] & bit})
        return {"sm_mhz": float(np.median([s[1] for s in timed])), "sm_max_mhz": self.max_sm, "reasons": seen,
                "samples": len(timed)}


def cpu_port_step(spec, p, t, w):
    """The oracle port of the reference's CPU path for the same step: serial per-image loss with autograd
    backward (tf.map_fn(parallel_iterations=1) + TF autodiff), decode, filter_boxes, apply_nms per image
    (run.py:59-74).  Returns nothing; timed by the caller."""
    import torch
    from oracle import ref_numpy as rn, ref_postprocess as rp, ref_torch as rt
    rt.loss_and_grad(t, p, spec, rectified_active=w["rect"] >= 0, dtype=torch.float32)
    dec = rn.decode(p, spec)
    for i in range(p.shape[0]):
        filt = rp.filter_boxes(dec[0][1][i], dec[0][2][i], dec[1][1][i], dec[1][2][i], dec[2][1][i], dec[2][2][i],
                               w["score_thresh"])
        rp.apply_nms(filt, w["nms_thresh"])


def time_cpu_port(spec, p, t, w, budget_s):
    """images/s of the oracle port on a bounded sample of the batch (about budget_s seconds of CPU work)."""
    import torch
    t0 = time.time()
    cpu_port_step(spec, p[:1], t[:1], w)
    one = max(time.time() - t0, 1e-3)
    s = int(max(1, min(p.shape[0], budget_s / one)))
    t0 = time.time()
    cpu_port_step(spec, p[:s], t[:s], w)
    dt = time.time() - t0
    return s / dt, s, torch.get_num_threads()


def run_reference(args, w, spec, rank, world):
    if rank != 0:
        return
    import torch
    # torchrun exports OMP_NUM_THREADS=1; the CPU arm gets every core this process may run on
    try:
        torch.set_num_threads(max(1, len(os.sched_getaffinity(0))))
    except Exception:
        pass
    p, t = synth(spec, w["batch"], w["m"], 0)
    cpu_port_step(spec, p[:1], t[:1], w)                       # import / first-touch warm-up
    t0 = time.time()
    cpu_port_step(spec, p[:1], t[:1], w)
    one = max(time.time() - t0, 1e-3)
    total_steps = args.steps + args.warmup
    s = int(max(1, min(w["batch"], (120.0 / total_steps) / one)))     # whole run within a few minutes
    for _ in range(args.warmup):
        cpu_port_step(spec, p[:s], t[:s], w)
    t0 = time.time()
    for _ in range(args.steps):
        cpu_port_step(spec, p[:s], t[:s], w)
    dt = time.time() - t0
    value = s * args.steps / dt
    cores = torch.get_num_threads()
    sample = "%d of %d images per step, fp32 torch-CPU restatement of the TF graph + NumPy post-process" % (s, w["batch"])
    print(json.dumps({
        "impl": "reference", "metric": "head loss fwd+bwd + decode+filter+NMS throughput", "value": value,
        "unit": "images/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "desc": w["desc"], "images_per_step": s},
        "cpu_baseline": {"value": value, "unit": "images/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "TensorFlow is not installable on this image (no network); the reference's TF loss/decoder "
                "cannot run, so the timed CPU arm is the oracle port (oracle/ref_torch.py, ref_numpy.py, "
                "ref_postprocess.py) on all host threads.",
    }), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg2_416_c80_b64", choices=sorted(WORKLOADS))
    ap.add_argument("--cpu-budget", type=float, default=15.0, help="seconds of CPU work for cpu_baseline")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--overlap", type=int, default=1, help="1: loss and post-process halves of a step on two streams")
    ap.add_argument("--priority", default="post", choices=["none", "post", "loss"],
                    help="which half of a step runs on a high-priority stream (two-stream mode)")
    ap.add_argument("--layout", default="packed", choices=["packed", "heads"],
                    help="packed: the reference's concatenated head tensor; heads: the three head tensors before "
                         "yolov3_detector.py:80-85 packs them (yv3_loss_fwd_bwd_heads / yv3_filter_heads)")
    ap.add_argument("--serialize-sweeps", type=int, default=-1,
                    help="1: the two HBM-bound sweeps (dense pass, score pass) never run at the same time; the small "
                         "kernels of either half run under the other half's sweep (default: on with the "
                         "packed layout, where several steps share a graph; measured 142 -> 140.5 us/step)")
    ap.add_argument("--fused", type=int, default=1,
                    help="1 (packed layout): loss fwd+bwd and the post-process filter of a step in one call that reads the "
                         "head tensor once (yv3_loss_filter_fused); 0: the two sweeps as separate calls")
    ap.add_argument("--fused-depth", type=int, default=3,
                    help="workspace pairs the fused schedule alternates between (the assignment of step k + depth follows "
                         "the fold of step k)")
    ap.add_argument("--graph-steps", type=int, default=16, help="steps captured per CUDA graph (N = 1)")
    ap.add_argument("--pipeline", type=int, default=1,
                    help="1: the two streams are joined once per graph / timed region, 0: after every step")
    ap.add_argument("--force-graph", action="store_true", help="skip the eager-vs-graph trial")
    ap.add_argument("--no-graph", action="store_true", help="time eager launches instead of a CUDA-graph replay")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    w = WORKLOADS[args.workload]
    spec = build_spec(w) if args.impl == "reference" else product_spec(w)
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.layout != "packed":
        args.fused = 0
    if args.fused:
        args.serialize_sweeps = 0
    if args.serialize_sweeps < 0:
        args.serialize_sweeps = 1 if (args.layout == "packed" and args.overlap and args.pipeline) else 0
    if args.impl == "reference":
        return run_reference(args, w, spec, rank, world)

    import torch
    import torch.distributed as dist
    import yolov3_tensorflow_b200 as yv3

    assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU path in the product)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    yv3.backend.set_epsilon(spec.eps)
    n, m = w["batch"], w["m"]
    p_np, t_np = synth(spec, n, m, rank)
    p_host = torch.from_numpy(p_np).pin_memory()
    t_host = torch.from_numpy(t_np).pin_memory()
    pd, td = p_host.to(dev), t_host.to(dev)
    loss = yv3.YOLOv3Loss(rectified_coord_num=w["rect"], **spec.loss_args)
    post = yv3.HeadPostProcessor(spec.head_grid_sizes, spec.class_num, spec.anchor_boxes)
    # a second workspace pair for the fused schedule (steps alternate; the C ABI is re-entrant per workspace)
    pairs = [(loss, post)]
    for _ in range(max(args.fused_depth, 2) - 1 if args.fused else 0):
        pairs.append((yv3.YOLOv3Loss(rectified_coord_num=w["rect"], **spec.loss_args),
                      yv3.HeadPostProcessor(spec.head_grid_sizes, spec.class_num, spec.anchor_boxes)))

    if args.fused:
        # the one-sweep kernel covers the packed, aligned case up to 262144 tiles per call; beyond it the call runs the two
        # sweeps itself, and the bench then schedules them as separate calls (--fused 0) and says so
        loss.loss_grad_and_filter(td, pd, post, w["score_thresh"])
        torch.cuda.synchronize()
        if not loss.last_fused:
            args.fused = 0
            pairs = pairs[:1]
            args.serialize_sweeps = 1 if (args.overlap and args.pipeline) else 0

    def set_auto(v):
        for lo, _ in pairs:
            lo.auto_all_reduce = v

    grad = torch.empty_like(pd)
    cand = torch.empty((n, post.capacity, 8), dtype=torch.float32, device=dev)
    counts = torch.empty((n, 3), dtype=torch.int32, device=dev)
    keep = torch.empty((n, post.capacity), dtype=torch.uint8, device=dev)

    post_stream = None
    if args.overlap:
        # the block scheduler serves the high-priority stream first whenever an SM slot frees up
        post_stream = torch.cuda.Stream(priority=-1 if args.priority == "post" else 0)
        aux_stream = torch.cuda.Stream(priority=-1 if args.priority == "post" else 0)       # fused schedule: assignment and fold
        if args.priority == "loss":
            torch.cuda.set_stream(torch.cuda.Stream(priority=-1))

    heads = hgrads = None
    if args.layout == "heads":
        # the same logits as three contiguous head tensors (N, H, W, B*L): what the head convolutions emit
        heads = []
        for h in range(3):
            gh, gw = spec.head_grid_sizes[h]
            s0, s1 = spec.chan_base[h], spec.chan_base[h] + spec.head_channels[h]
            heads.append(pd[..., s0:s1].reshape(n, gh, gw, -1).contiguous())
        hgrads = [torch.empty_like(x) for x in heads]

    def half_loss(events=None, sweep_after=None, sweep_done=None, out=None):
        if sweep_after is not None or sweep_done is not None:
            total, _ = loss.loss_and_grad(td, pd, grad_out=grad, sweep_after=sweep_after, sweep_done=sweep_done, out=out)
            return total
        if heads is not None and events is None:
            total, _ = loss.loss_and_grad_heads(td, heads, grads_out=hgrads)
            return total
        total, _ = loss.loss_and_grad(td, pd, events=events, grad_out=grad, out=out)
        return total

    def half_post():
        if heads is not None:
            post.filter_heads(heads, w["score_thresh"], out=(cand, counts))
        else:
            post.filter(pd, w["score_thresh"], out=(cand, counts))
        post.nms(cand, counts, w["nms_thresh"], out=keep)

    def fused_steps(count, events=None, outs=None):
        """`count` complete steps with the one-sweep call (yv3_loss_filter_fused issued phase by phase), two (loss,
        post-process) workspace pairs alternating by step: the HBM-bound sweeps run back to back on the current stream,
        the small kernels under them on two other streams (see below).  Every step does all of its work."""
        thr = w["score_thresh"]
        depth = len(pairs)
        cur = torch.cuda.current_stream()
        if post_stream is None or events is not None:
            for k in range(count):
                total, _, _, _ = loss.loss_grad_and_filter(td, pd, post, thr, grad_out=grad,
                                                           out=None if outs is None else outs[k],
                                                           filter_out=(cand, counts), events=events)
                post.nms(cand, counts, w["nms_thresh"], out=keep)
            return total

        def part(k, phases):
            lo, po = pairs[k % depth]
            return lo.loss_grad_and_filter(td, pd, po, thr, grad_out=grad, out=None if outs is None else outs[k],
                                           filter_out=(cand, counts), phases=phases)[0]

        def assign(k):
            with torch.cuda.stream(aux_stream):
                part(k, 1)
                e = torch.cuda.Event()
                e.record(aux_stream)
            return e

        # three streams, `depth` workspace pairs used round-robin.  cur: the sweeps, back to back.  aux: fold(k) then
        # assign(k + depth) -- the assignment rewrites the loss workspace step k used, so it follows that step's sweep and
        # fold.  post: gather(k), sort, NMS(k); sweep(k + depth) rewrites the staging rows gather(k) reads, so it waits for
        # that chain (long finished by then).  depth 3 instead of 2 gives the fold -> assignment chain a whole extra sweep
        # of slack before sweep(k + depth) needs it (measured 104.8 -> 99.6 us/step at cfg2; depth 4: 99.2).
        aux_stream.wait_stream(cur)
        post_stream.wait_stream(cur)
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
            with torch.cuda.stream(post_stream):
                post_stream.wait_event(swept)
                part(k, 8)
                pairs[b][1].nms(cand, counts, w["nms_thresh"], out=keep)
                g_done[b] = torch.cuda.Event()
                g_done[b].record(post_stream)
            with torch.cuda.stream(aux_stream):
                aux_stream.wait_event(swept)
                total = part(k, 4)
            if k + depth < count:
                a_done[b] = assign(k + depth)
        cur.wait_stream(post_stream)
        cur.wait_stream(aux_stream)
        return total

    def steps(count, events=None, outs=None):
        """`count` complete steps.  The training half (assignment, dense pass) and the inference half (filter, sort,
        NMS) of a step share only their input, so they run on two streams.  --pipeline 1: the streams are joined once
        per call (consecutive steps of either half queue behind each other, the latency-bound tail of one step's
        post-process runs under the next step's sweeps); --pipeline 0: joined after every step."""
        if args.fused:
            return fused_steps(count, events, outs)
        if post_stream is None or events is not None:
            for k in range(count):
                total = half_loss(events)
                half_post()
            return total
        cur = torch.cuda.current_stream()
        post_stream.wait_stream(cur)
        dense_done = None
        for k in range(count):
            if args.serialize_sweeps and heads is None:
                with torch.cuda.stream(post_stream):
                    if dense_done is not None:
                        post_stream.wait_event(dense_done)          # the score sweep waits for the previous dense sweep
                    filt_done = torch.cuda.Event()
                    post.filter(pd, w["score_thresh"], out=(cand, counts), sweep_done=filt_done)   # recorded after the sweep
                    post.nms(cand, counts, w["nms_thresh"], out=keep)
                dense_done = torch.cuda.Event()
                total = half_loss(sweep_after=filt_done, sweep_done=dense_done,      # the dense sweep waits for the score sweep
                                  out=None if outs is None else outs[k])
                continue
            with torch.cuda.stream(post_stream):
                half_post()
            total = half_loss(out=None if (outs is None or heads is not None) else outs[k])
            if not args.pipeline:
                cur.wait_stream(post_stream)
                if k + 1 < count:
                    post_stream.wait_stream(cur)
        if args.pipeline:
            cur.wait_stream(post_stream)
        return total

    def step(events=None):
        return steps(1, events)

    def fence():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # owned-label count for the algorithmic byte figure: R = sum of `owned` flags
    table, _ = loss.assign(td, pd)
    owned = int(table[..., 3].sum().item())
    loss.check_status()

    sampler = ClockSampler(local_rank)
    sampler.start()
    sampler.mark = "warmup"
    for _ in range(args.warmup):
        step()
    fence()
    # per-kernel view: eager launches with CUDA events around the dense sweep and the post-process (same K steps)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    post_ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    sampler.mark = "kernels"
    for k in range(args.steps):
        step(ev[k])
        post_ev[k].record()
    fence()
    dense_ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    post_ms = float(np.mean([ev[k][1].elapsed_time(post_ev[k]) for k in range(args.steps)]))

    # the timed region: the whole step (assignment, dense pass, all-reduce at N > 1, filter, NMS) captured once in a
    # CUDA graph and replayed K times -- the launch-bound host loop is out of the way, the device work is identical
    graph, mode = None, "eager"
    # several steps per graph (no collective inside at N = 1): consecutive steps overlap like eager launches do
    # N > 1: the NCCL call stays outside the graph (a collective inside the capture hung in round 1); the steps of a
    # graph write their 20 loss floats into the rows of one buffer, which is all-reduced once after the replay
    unroll = args.graph_steps if (world == 1 or args.layout == "packed") else 1
    loss_rows = torch.zeros((unroll, 20), dtype=torch.float32, device=dev) if world > 1 else None
    if not args.no_graph:
        try:
            set_auto(False)                           # the collective stays outside the graph
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                step()
            torch.cuda.current_stream().wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                steps(unroll, outs=loss_rows)
            mode = "cuda-graph replay, %d step(s) per graph" % unroll + (
                " + one eager NCCL all-reduce of the graph's %d x 20 loss floats" % unroll if world > 1 else "")
        except Exception as e:
            graph, mode = None, "eager (graph capture failed: %s)" % type(e).__name__
            torch.cuda.synchronize()
    if graph is None:
        set_auto(True)

    def run(count):
        """exactly `count` steps"""
        if graph is None:
            steps(count)
            return
        for _ in range(count // unroll):
            graph.replay()
            if world > 1:
                dist.all_reduce(loss_rows)
        rest = count % unroll
        if rest:
            set_auto(True)
            steps(rest)
            set_auto(False)
    run(3 * unroll)
    fence()
    if graph is not None and not args.force_graph:
        # launch mode by trial: the host may launch eagerly fast enough for consecutive steps to overlap better than
        # graph replays do; 20 steps each, the faster one is used for the timed region
        def trial(fn):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            fence()
            a.record()
            fn()
            b.record()
            fence()
            return a.elapsed_time(b)

        def eager20():
            set_auto(True)
            steps(20)
            set_auto(False)

        t_graph = trial(lambda: run(20 - 20 % unroll))
        t_eager = trial(eager20)
        verdict = torch.tensor([t_graph, t_eager], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(verdict, op=dist.ReduceOp.MAX)
        if float(verdict[1]) < float(verdict[0]):
            graph, mode = None, "eager launches (faster than graph replay in a 20-step trial: %.1f vs %.1f us/step)" % (
                1e3 * float(verdict[1]) / 20, 1e3 * float(verdict[0]) / (20 - 20 % unroll))
            set_auto(True)
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.mark = "timed"
    start.record()
    run(args.steps)
    stop.record()
    fence()
    sampler.mark = "after"
    total_ms = start.elapsed_time(stop)
    for lo, po in pairs:
        lo.check_status()
        po.check_status()
    cand_per_image = float(counts.sum().item()) / n
    kept_per_image = float(sum(int(keep[i, :int(counts[i].sum())].sum()) for i in range(min(n, 8)))) / min(n, 8)

    # ---- end to end through the public API with host buffers -------------------------------------------

    set_auto(True)

    # Double-buffered: the host->device copy of step k + 1 (copy stream) runs while step k computes; the caller reads
    # step k - 1's loss and detections from pinned host memory once that step's device->host copies have landed.
    # Every step's inputs cross PCIe inside the timed region and every step's results are read on the host.
    copy_stream = torch.cuda.Stream()
    pbuf = [torch.empty_like(pd) for _ in range(2)]
    tbuf = [torch.empty_like(td) for _ in range(2)]
    out_hosts = [torch.empty(20, dtype=torch.float32).pin_memory() for _ in range(2)]
    counts_hosts = [torch.empty((n, 3), dtype=torch.int32).pin_memory() for _ in range(2)]
    keep_hosts = [torch.empty((n, post.capacity), dtype=torch.uint8).pin_memory() for _ in range(2)]

    def e2e_run(count):
        main = torch.cuda.current_stream()
        done = [None, None]
        read = 0.0
        for k in range(count):
            b = k & 1
            with torch.cuda.stream(copy_stream):
                if done[b] is not None:
                    copy_stream.wait_event(done[b])                   # step k - 2 is finished with this buffer pair
                pbuf[b].copy_(p_host, non_blocking=True)              # pinned host -> device, every step
                tbuf[b].copy_(t_host, non_blocking=True)
                ready = torch.cuda.Event()
                ready.record(copy_stream)
            main.wait_event(ready)
            if args.fused:
                total, _, _, _ = loss.loss_grad_and_filter(tbuf[b], pbuf[b], post, w["score_thresh"], grad_out=grad,
                                                           filter_out=(cand, counts))
            else:
                total, _ = loss.loss_and_grad(tbuf[b], pbuf[b], grad_out=grad)
                post.filter(pbuf[b], w["score_thresh"], out=(cand, counts))
            post.nms(cand, counts, w["nms_thresh"], out=keep)
            out_hosts[b][:18].copy_(loss.loss_detail.reshape(-1), non_blocking=True)
            out_hosts[b][18:19].copy_(total.reshape(1), non_blocking=True)
            counts_hosts[b].copy_(counts, non_blocking=True)
            keep_hosts[b].copy_(keep, non_blocking=True)
            done[b] = torch.cuda.Event()
            done[b].record(main)
            if k > 0:
                done[b ^ 1].synchronize()                             # the caller reads step k - 1's results
                read += float(out_hosts[b ^ 1][18]) + float(counts_hosts[b ^ 1][0, 0])
        done[(count - 1) & 1].synchronize()
        read += float(out_hosts[(count - 1) & 1][18])
        return read

    sampler.mark = "e2e"
    e2e_run(3)
    fence()
    e2e_steps = args.steps
    s2, t2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s2.record()
    e2e_run(e2e_steps)
    t2.record()
    fence()
    e2e_ms = s2.elapsed_time(t2)
    sampler.stop_flag = True
    sampler.join(timeout=2)

    times = torch.tensor([total_ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms = float(times[0]), float(times[1])

    if rank == 0:
        A, L, C = spec.anchors_per_image, spec.box_len, spec.class_num
        alg_per_image = 4 * A * 5 + 4 * A * L + 4 * (owned / n) * C + 4 * m * 5
        kernel_name = "loss_dense_kernel"
        if args.fused:
            # the fused sweep does the loss sweep's AND the score sweep's work: SURVEY 8(d)'s figures for the two added
            # up with the tensor read counted once -- 4 A L in (every class logit feeds max / arg-max), 4 A L out
            # (the dense gradient), the labels; the survivors' staging rows are noise
            alg_per_image = 4 * A * L + 4 * A * L + 4 * m * 5
            kernel_name = "fused_sweep_kernel"
            assert all(lo.last_fused for lo, _ in pairs), "the fused call fell back to the two sweeps"
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        else:
            peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
        achieved = alg_per_image * n / (dense_ms * 1e-3) / 1e9
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            traffic = json.load(open(tp)).get(args.workload, {}).get(kernel_name + "_dram_bytes_per_launch")
        line = {
            "metric": "head loss fwd+bwd + decode+filter+NMS throughput",
            "value": world * n * args.steps / (total_ms * 1e-3), "unit": "images/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "desc": w["desc"], "per_gpu_batch": n, "global_batch": n * world,
                       "label_slots": m, "score_thresh": w["score_thresh"], "nms_thresh": w["nms_thresh"],
                       "parallelism": "batch-sharded dp%d, the 20 loss floats of every step all-reduced over NCCL "
                                      "(one call per graph replay / per eager step)" % world,
                       "l2": "inputs larger than L2: predicts %.0f MB + gradient %.0f MB per GPU vs 126 MB L2" % (
                           pd.numel() * 4 / 1e6, grad.numel() * 4 / 1e6),
                       "candidates_per_image": cand_per_image, "kept_per_image": kept_per_image,
                       "layout": args.layout, "fused_sweep": bool(args.fused), "serialize_sweeps": bool(args.serialize_sweeps), "launch": mode, "streams": (3 if args.fused else 2) if args.overlap else 1,
                       "high_priority_stream": args.priority if args.overlap else None,
                       "fused_schedule": "three streams: sweeps back to back; fold of step k and assignment of step k+depth; "
                                         "gather/sort/NMS of step k (the last two at high priority); %d workspaces alternating" % len(pairs) if (args.fused and args.overlap) else None,
                       "stream_join": ("once per %d-step graph" % unroll if graph is not None else "once per timed region")
                       if (args.overlap and args.pipeline) else "every step"},
            "stage_ms": {kernel_name: dense_ms, ("fold_gather_sort_nms" if args.fused else "filter_plus_nms"): post_ms,
                         "step_minus_stages": total_ms / args.steps - dense_ms - post_ms,
                         "note": "stage times from eager single-stream launches with events; ms_per_step from the timed "
                                 "region (step_minus_stages < 0: the two halves overlap on two streams; it also holds "
                                 "the assignment and fold kernels)"},
            "roofline": {"kernel": kernel_name, "bound": "hbm", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": alg_per_image * n, "avg_launch_ms": dense_ms,
                         "dense_2x4AL_bytes_per_launch_for_context": 8 * A * L * n},
            "e2e": {"value": world * n * e2e_steps / (e2e_ms * 1e-3), "unit": "images/s",
                    "h2d_bytes_per_step": int(p_host.numel() * 4 + t_host.numel() * 4),
                    "d2h_bytes_per_step": int(19 * 4 + counts_hosts[0].numel() * 4 + keep_hosts[0].numel()),
                    "pipelining": "double-buffered: the copy of step k+1 overlaps the compute of step k; results of "
                                  "step k-1 are read on the host every step",
                    "ms_per_step": e2e_ms / e2e_steps, "steps": e2e_steps},
            # assign, dense sweep, fold, score, gather, sort, nms x2 (the sweeps run once per head tensor in the heads layout)
            # fused: assign, fused sweep, fold, gather, sort, nms x2
            "gpu_launches": (12 if args.layout == "heads" else 7 if args.fused else 8) * args.steps,
            "clocks": sampler.summary(),
        }
        if world == 1 and not args.no_cpu_baseline:
            v, s, cores = time_cpu_port(build_spec(w), p_np, t_np, w, args.cpu_budget)       # the oracle enters here only
            line["cpu_baseline"] = {"value": v, "unit": "images/s", "cores": cores, "kind": "port",
                                    "sample": "%d of %d images, one pass (oracle port: torch-CPU fp32 loss + autograd, "
                                              "NumPy decode/filter, Python NMS as run.py runs it)" % (s, n)}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
