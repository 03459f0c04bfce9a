"""Benchmark of the detection-head hot path (BASELINE.json): fused head loss fwd+bwd and decode+filter+NMS.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A step = one pass of the hot path over one synthetic batch: assignment + fused loss forward/backward (gradient written)
and fused decode + score threshold + class-aware NMS of the same head tensors, issued through
``yolov3_tensorflow_b200.FusedPipeline`` (the schedule tests/test_gpu_pipeline.py checks bit for bit).
One JSON line on stdout (rank 0).  See DESIGN.md "Measurement" for every field.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOADS = {
    # BASELINE.json configs[1]: the configuration the metric is quoted on at N=1 (weak scaling: 64 images per GPU)
    "cfg2_416_c80_b64": dict(size=(416, 416), class_num=80, batch=64, m=16, anchors="coco", iou_thresh=0.8,
                             focal=False, rect=-1, score_thresh=0.1, nms_thresh=0.45,
                             desc="416x416 80-class head (10647 boxes/img), batch 64 per GPU, loss fwd+bwd + "
                                  "decode+filter+NMS"),
    # configs[2] per-GPU shard at 8 GPUs (weak form)
    "cfg3_608_c80_b64": dict(size=(608, 608), class_num=80, batch=64, m=16, anchors="coco", iou_thresh=0.8,
                             focal=False, rect=-1, score_thresh=0.1, nms_thresh=0.45,
                             desc="608x608 80-class head (22743 boxes/img), batch 64 per GPU"),
    # configs[2] as written: global batch 512 SHARDED over the GPUs (512 / 256 / 128 / 64 images per rank)
    "cfg3_608_c80_g512": dict(size=(608, 608), class_num=80, global_batch=512, m=16, anchors="coco", iou_thresh=0.8,
                              focal=False, rect=-1, score_thresh=0.1, nms_thresh=0.45,
                              desc="608x608 80-class head (22743 boxes/img), GLOBAL batch 512 sharded over the GPUs"),
    # configs[3]: post-process stress
    "cfg4_416_c80_b256_dense": dict(size=(416, 416), class_num=80, batch=256, m=16, anchors="coco", iou_thresh=0.8,
                                    focal=False, rect=-1, score_thresh=0.001, nms_thresh=0.45,
                                    desc="416x416 batch 256, score thresh 0.001, per-class NMS IoU 0.45 (dense)"),
    # configs[4] as written: a LOSS configuration -- single class (class_num 0), focal + prior-box correction on, 512x512,
    # global batch 128 over the GPUs (16 per rank at 8); a score threshold that leaves the usual handful of candidates
    "cfg5_512_c0_g128": dict(size=(512, 512), class_num=0, global_batch=128, m=16, anchors="shipped3", iou_thresh=0.8,
                             focal=True, rect=1 << 30, score_thresh=0.95, nms_thresh=0.4,
                             desc="512x512 single-class head (class_num 0), focal + prior-box correction on, GLOBAL batch 128 "
                                  "sharded over the GPUs"),
    # the one-class NMS stress round 1 ran under the name cfg5 (score threshold 0.5 on N(0,1) logits: ~8000 candidates of
    # ONE class per image) -- a post-process stress, not BASELINE configs[4]
    "cfg5_nms_stress_b16": dict(size=(512, 512), class_num=0, batch=16, m=16, anchors="shipped3", iou_thresh=0.8,
                                focal=True, rect=1 << 30, score_thresh=0.5, nms_thresh=0.4,
                                desc="512x512 class_num 0, batch 16, score thresh 0.5: one class of ~8000 candidates per image"),
}
EXTRA_WORKLOADS = ("cfg3_608_c80_g512", "cfg5_512_c0_g128")     # short entries appended to every line (BASELINE configs[2], [4])


def per_rank_batch(w, world):
    if "global_batch" in w:
        if w["global_batch"] % world:
            raise ValueError("global batch %d does not divide over %d ranks" % (w["global_batch"], world))
        return w["global_batch"] // world
    return w["batch"]


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
    rank = int(os.environ.get("YV3_BENCH_SEED_RANK", rank))      # developer switch: another rank's data on this GPU
    h32, w32 = spec.head_grid_sizes[2]
    return (ps.make_predicts(n, h32 * 32, w32 * 32, spec.box_num, spec.class_num, seed=1234 + rank),
            ps.make_targets(n, spec.class_num, m, seed=4321 + rank, min_valid=1))


def config_block(name, w, n, world, extra=None):
    """the ``config`` object: the same keys in the product arm and the reference arm"""
    c = {"workload": name, "desc": w["desc"], "per_gpu_batch": n, "global_batch": n * world, "label_slots": w["m"],
         "score_thresh": w["score_thresh"], "nms_thresh": w["nms_thresh"]}
    c.update(extra or {})
    return c


def git_sha():
    try:
        return subprocess.run(["git", "rev-parse", "--short", "HEAD"], cwd=ROOT, capture_output=True, text=True).stdout.strip() or None
    except Exception:
        return None


def bind_to_gpu_numa_node(index):
    """N > 1 only: run this rank (and so first-touch its pinned host buffers) on the CPUs of the NUMA node its GPU hangs
    off -- the e2e leg copies 232 MB per step and rank from pinned host memory, and with all ranks on one node's memory the
    copies share that node's controllers.  Best effort: returns what was done, or why not."""
    try:
        import pynvml
        pynvml.nvmlInit()
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(index)).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        dom, rest = bus.split(":", 1)
        path = "/sys/bus/pci/devices/%s:%s/numa_node" % (dom[-4:].lower(), rest.lower())
        node = int(open(path).read().strip())
        if node < 0:
            return {"bound": False, "why": "the GPU reports no NUMA node"}
        cpus = set()
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return {"bound": False, "why": "no allowed CPU on node %d" % node}
        os.sched_setaffinity(0, cpus)
        return {"bound": True, "node": node, "cpus": len(cpus)}
    except Exception as e:                                          # noqa: BLE001
        return {"bound": False, "why": "%s: %s" % (type(e).__name__, str(e)[:120])}


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
                try:
                    mem = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_MEM)
                except Exception:
                    mem = None
                self.samples.append((time.time(), sm, reasons, self.mark, mem))
            except Exception:
                pass
            time.sleep(0.005)

    def summary(self):
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_sm, "reasons": ["unavailable"]}
        nv = self.nv
        timed = [s for s in self.samples if s[3] == "timed"] or [s for s in self.samples if s[3] is not None] \
            or self.samples
        names = {
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
            "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
            "hw_power_brake": getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80),
        }
        seen = sorted({k for s in timed for k, bit in names.items() if s[2] & bit})
        mems = [s[4] for s in timed if s[4] is not None]
        return {"sm_mhz": float(np.median([s[1] for s in timed])), "sm_max_mhz": self.max_sm, "reasons": seen,
                "samples": len(timed), "mem_mhz": float(np.median(mems)) if mems else None}


# ---- CPU legs: the only place bench.py executes oracle/ ---------------------------------------------------------------
def cpu_port_step(spec, p, t, w):
    """The oracle port of the reference's CPU path for the same step: serial per-image loss with autograd backward
    (tf.map_fn(parallel_iterations=1) + TF autodiff), decode, filter_boxes, apply_nms per image (run.py:59-74)."""
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
    n = per_rank_batch(w, world)
    p, t = synth(spec, n, w["m"], 0)
    cpu_port_step(spec, p[:1], t[:1], w)                       # import / first-touch warm-up
    t0 = time.time()
    cpu_port_step(spec, p[:1], t[:1], w)
    one = max(time.time() - t0, 1e-3)
    total_steps = args.steps + args.warmup
    s = int(max(1, min(n, (120.0 / total_steps) / one)))     # whole run within a few minutes
    for _ in range(args.warmup):
        cpu_port_step(spec, p[:s], t[:s], w)
    t0 = time.time()
    for _ in range(args.steps):
        cpu_port_step(spec, p[:s], t[:s], w)
    dt = time.time() - t0
    value = s * args.steps / dt
    cores = torch.get_num_threads()
    sample = "%d of %d images per step, fp32 torch-CPU restatement of the TF graph + NumPy post-process" % (s, n)
    print(json.dumps({
        "impl": "reference", "metric": "head loss fwd+bwd + decode+filter+NMS throughput", "value": value,
        "unit": "images/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": config_block(args.workload, w, n, world, {"images_per_step_sampled": s}),
        "cpu_baseline": {"value": value, "unit": "images/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "TensorFlow is not installable on this image (no network); the reference's TF loss/decoder "
                "cannot run on the GPU box, so the timed CPU arm is the oracle port (oracle/ref_torch.py, ref_numpy.py, "
                "ref_postprocess.py), which tests/test_oracle_pinned.py holds bit for bit (fp32) to the reference's own "
                "code run under the torch-backed tensorflow stand-in in the build container.",
    }), flush=True)


# ---- the product arm ----------------------------------------------------------------------------------------------------
class Arm(object):
    """One workload resident on this rank's GPU: inputs, the pipeline, its output buffers, and the measurements."""

    def __init__(self, name, args, rank, world, dev, reducer, pin=True):
        import torch
        import yolov3_tensorflow_b200 as yv3
        self.torch, self.yv3 = torch, yv3
        self.name, self.w, self.args = name, WORKLOADS[name], args
        self.rank, self.world, self.dev = rank, world, dev
        w = self.w
        self.spec = product_spec(w)
        yv3.backend.set_epsilon(self.spec.eps)
        self.n, self.m = per_rank_batch(w, world), w["m"]
        self.p_np, self.t_np = synth(self.spec, self.n, self.m, rank)
        self.p_host, self.t_host = torch.from_numpy(self.p_np), torch.from_numpy(self.t_np)
        if pin:                                                    # the e2e loop copies from pinned host memory every step
            self.p_host, self.t_host = self.p_host.pin_memory(), self.t_host.pin_memory()
        self.pd, self.td = self.p_host.to(dev), self.t_host.to(dev)
        self.pipe = yv3.FusedPipeline(self.spec.loss_args, w["score_thresh"], w["nms_thresh"], depth=max(args.fused_depth, 1),
                                      rectified_coord_num=w["rect"], reducer=reducer)
        self.loss, self.post = self.pipe.pairs[0]
        self.grad = torch.empty_like(self.pd)
        cap = self.pipe.capacity
        self.cand = torch.empty((self.n, cap, 8), dtype=torch.float32, device=dev)
        self.counts = torch.empty((self.n, 3), dtype=torch.int32, device=dev)
        self.keep = torch.empty((self.n, cap), dtype=torch.uint8, device=dev)
        self.unroll = max(1, args.graph_steps)
        self.rows = torch.zeros((self.unroll, 20), dtype=torch.float32, device=dev)
        self.graphs = {}

    # -- issue `count` steps (count <= unroll): the pipeline, or the separate calls one after the other
    def issue(self, count):
        if self.args.fused:
            self.pipe.run(self.td, self.pd, self.grad, self.rows, self.cand, self.counts, self.keep, count=count)
            if self.pipe.reducer is None and self.world > 1 and not self.capturing:
                self.pipe.all_reduce_rows(self.rows)
        else:
            self.loss.auto_all_reduce = not self.capturing
            for k in range(count):
                self.loss.loss_and_grad(self.td, self.pd, grad_out=self.grad, out=self.rows[k])
                self.post.filter(self.pd, self.w["score_thresh"], out=(self.cand, self.counts))
                self.post.nms(self.cand, self.counts, self.w["nms_thresh"], out=self.keep)
    capturing = False

    def graph_for(self, count):
        """a CUDA graph of `count` steps (cached); None when capture is off or failed"""
        torch = self.torch
        if self.args.no_graph:
            return None
        if count not in self.graphs:
            try:
                self.issue(count)                                  # warm-up outside the capture
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                self.capturing = True
                with torch.cuda.graph(g):
                    self.issue(count)
                self.graphs[count] = g
            except Exception as e:                                 # noqa: BLE001
                self.graphs[count] = None
                self.capture_error = type(e).__name__
                torch.cuda.synchronize()
            finally:
                self.capturing = False
        return self.graphs[count]

    def run(self, count):
        """exactly `count` steps: full graphs, then ONE graph of the remainder (no eager tail)"""
        done = 0
        while done < count:
            c = min(self.unroll, count - done)
            g = self.graph_for(c)
            if g is None:
                self.issue(c)
            else:
                g.replay()
                if self.pipe.reducer is None and self.world > 1:
                    self.pipe.all_reduce_rows(self.rows)           # NCCL stays outside the graph: once per replay
            done += c

    def launch_mode(self):
        if self.graphs.get(self.unroll) is None:
            return "eager launches" + (" (graph capture failed: %s)" % getattr(self, "capture_error", "?") if not self.args.no_graph else "")
        red = ("the peer-memory all-reduce kernel of every step's 20 loss floats inside the graph" if self.pipe.reducer is not None
               else "one eager NCCL all-reduce of the graph's %d x 20 loss floats per replay" % self.unroll) if self.world > 1 else None
        return "cuda-graph replay, %d step(s) per graph, the remainder as one smaller graph" % self.unroll + (" + " + red if red else "")

    def fence(self):
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier()
        self.torch.cuda.synchronize()

    def time_steps(self, steps, min_ms=100.0):
        """CUDA events around `steps` steps repeated until the region is at least min_ms long.
        Returns (ms per step, steps timed)."""
        torch = self.torch
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.run(steps)                                            # graphs built, caches in their steady state
        self.fence()
        a.record()
        self.run(steps)
        b.record()
        self.fence()
        est = max(a.elapsed_time(b), 1e-3)
        reps = int(min(max(1, np.ceil(min_ms / est)), 2000))
        if self.world > 1:                                         # every rank must issue the same number of collectives
            import torch.distributed as dist
            r = torch.tensor([reps], device=self.dev)
            dist.all_reduce(r, op=dist.ReduceOp.MAX)
            reps = int(r)
        self.fence()
        a.record()
        for _ in range(reps):
            self.run(steps)
        b.record()
        self.fence()
        ms = a.elapsed_time(b)
        self.per_rank_ms = None
        if self.world > 1:
            import torch.distributed as dist
            t = torch.tensor([ms], dtype=torch.float64, device=self.dev)
            allt = [torch.zeros_like(t) for _ in range(self.world)]
            dist.all_gather(allt, t)
            self.per_rank_ms = [float(x) / (reps * steps) for x in allt]     # every rank's own clock around the same region
            ms = max(float(x) for x in allt)
        return ms / (reps * steps), reps * steps

    def time_sweep(self, reps):
        """the dominant kernel alone: CUDA events on the launching stream around the sweep of an eager step"""
        torch = self.torch
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
        for k in range(reps + 2):
            e = ev[max(k - 2, 0)]
            if self.args.fused:
                self.loss.loss_grad_and_filter(self.td, self.pd, self.post, self.w["score_thresh"], grad_out=self.grad,
                                               filter_out=(self.cand, self.counts), events=e, rect_override=self.w["rect"] >= 0)
            else:
                self.loss.loss_and_grad(self.td, self.pd, grad_out=self.grad, events=e)
            self.post.nms(self.cand, self.counts, self.w["nms_thresh"], out=self.keep)
        torch.cuda.synchronize()
        return float(np.mean([x.elapsed_time(y) for x, y in ev]))

    def alg_bytes(self, owned=0):
        """SURVEY 8(d) algorithmic bytes of one launch of the dominant kernel"""
        A, L, C, n, m = self.spec.anchors_per_image, self.spec.box_len, self.spec.class_num, self.n, self.m
        if self.args.fused:
            return n * (4 * A * L + 4 * A * L + 4 * m * 5), "fused_sweep_kernel"
        return n * (4 * A * 5 + 4 * A * L + 4 * m * 5) + 4 * owned * C, "loss_dense_kernel"

    # -- end to end through the public API with HOST buffers ---------------------------------------------------------------
    def time_e2e(self, steps):
        """Double-buffered: the host->device copy of step k + 1 (copy stream) runs while step k computes; the caller reads
        step k - 1's loss AND detections from pinned host memory once that step's device->host copies have landed.  Every
        step's inputs cross PCIe inside the timed region; the results read back are the 20 loss floats, the per-head counts,
        the keep flags, the candidate rows and the rows yv3_collect_kept returns (resized, float64).  The gradient stays on
        the device (it feeds the backward of the model there)."""
        torch = self.torch
        n, w, post, loss = self.n, self.w, self.post, self.loss
        size = (float(w["size"][1]), float(w["size"][0]))
        copy_stream = torch.cuda.Stream()
        pbuf = [torch.empty_like(self.pd) for _ in range(2)]
        tbuf = [torch.empty_like(self.td) for _ in range(2)]
        # rows a step returns: bounded by what the workload's threshold lets through (measured once), not by capacity
        kmax = int(self.counts.sum(dim=1).max().item())
        kcap = max(64, min(post.capacity, int(kmax * 1.25) + 64))
        out_h = [torch.empty(20, dtype=torch.float32).pin_memory() for _ in range(2)]
        counts_h = [torch.empty((n, 3), dtype=torch.int32).pin_memory() for _ in range(2)]
        kcounts_h = [torch.empty((n, 3), dtype=torch.int32).pin_memory() for _ in range(2)]
        keep_h = [torch.empty((n, kcap), dtype=torch.uint8).pin_memory() for _ in range(2)]
        cand_h = [torch.empty((n, kcap, 8), dtype=torch.float32).pin_memory() for _ in range(2)]
        rows_h = [torch.empty((n, kcap, 9), dtype=torch.float64).pin_memory() for _ in range(2)]
        row = torch.empty(20, dtype=torch.float32, device=self.dev)
        loss.auto_all_reduce = True
        self.e2e_d2h = int(20 * 4 + 2 * n * 3 * 4 + n * kcap * (1 + 32 + 72))
        self.e2e_h2d = int(self.p_host.numel() * 4 + self.t_host.numel() * 4)

        def run(count):
            main = torch.cuda.current_stream()
            done = [None, None]
            read = 0.0
            for k in range(count):
                b = k & 1
                with torch.cuda.stream(copy_stream):
                    if done[b] is not None:
                        copy_stream.wait_event(done[b])               # step k - 2 is finished with this buffer pair
                    pbuf[b].copy_(self.p_host, non_blocking=True)     # pinned host -> device, every step
                    tbuf[b].copy_(self.t_host, non_blocking=True)
                    ready = torch.cuda.Event()
                    ready.record(copy_stream)
                main.wait_event(ready)
                if self.args.fused:
                    loss.loss_grad_and_filter(tbuf[b], pbuf[b], post, w["score_thresh"], grad_out=self.grad, out=row,
                                              filter_out=(self.cand, self.counts), rect_override=w["rect"] >= 0)
                    if self.world > 1:
                        loss._all_reduce(row)
                else:
                    loss.loss_and_grad(tbuf[b], pbuf[b], grad_out=self.grad, out=row)
                    post.filter(pbuf[b], w["score_thresh"], out=(self.cand, self.counts))
                post.nms(self.cand, self.counts, w["nms_thresh"], out=self.keep)
                krows, kcounts = post.collect(dict(cand=self.cand, counts=self.counts, keep=self.keep), size)
                out_h[b].copy_(row, non_blocking=True)
                counts_h[b].copy_(self.counts, non_blocking=True)
                kcounts_h[b].copy_(kcounts, non_blocking=True)
                keep_h[b].copy_(self.keep[:, :kcap], non_blocking=True)
                cand_h[b].copy_(self.cand[:, :kcap], non_blocking=True)
                rows_h[b].copy_(krows[:, :kcap], non_blocking=True)
                done[b] = torch.cuda.Event()
                done[b].record(main)
                if k > 0:
                    done[b ^ 1].synchronize()                         # the caller reads step k - 1's results
                    read += float(out_h[b ^ 1][18]) + float(counts_h[b ^ 1][0, 0]) + float(rows_h[b ^ 1][0, 0, 7])
            done[(count - 1) & 1].synchronize()
            return read + float(out_h[(count - 1) & 1][18])

        run(3)
        self.fence()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        run(steps)
        b.record()
        self.fence()
        ms = a.elapsed_time(b)
        if self.world > 1:
            import torch.distributed as dist
            t = torch.tensor([ms], dtype=torch.float64, device=self.dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        return ms / steps

    def h2d_probe(self, reps=8):
        """the platform ceiling of the e2e number: bare pinned-host -> device copies of this rank's inputs, every rank
        copying at the same time; GB/s of this rank (max-time over ranks for the aggregate)"""
        torch = self.torch
        dst = torch.empty_like(self.pd)
        for _ in range(2):
            dst.copy_(self.p_host, non_blocking=True)
        self.fence()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            dst.copy_(self.p_host, non_blocking=True)
        b.record()
        self.fence()
        ms = a.elapsed_time(b)
        if self.world > 1:
            import torch.distributed as dist
            t = torch.tensor([ms], dtype=torch.float64, device=self.dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        nbytes = self.p_host.numel() * 4
        gbs = nbytes * reps / (ms * 1e-3) / 1e9
        return {"per_rank_gbs": gbs, "aggregate_gbs": gbs * self.world, "bytes_per_copy": int(nbytes),
                "ms_per_copy": ms / reps, "note": "bare pinned->device copy of one step's logits, all ranks at once, max over ranks"}

    # -- the stages taken alone (SURVEY 8(d) figures), CUDA events around eager launches ------------------------------------
    def stage_rooflines(self, peak, reps=20):
        torch, yv3 = self.torch, self.yv3
        A, L, C, n, m = self.spec.anchors_per_image, self.spec.box_len, self.spec.class_num, self.n, self.m
        table, _ = self.loss.assign(self.td, self.pd)
        owned = int(table[..., 3].sum().item())
        out = {}

        def timed(fn):
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
            for k in range(reps + 2):
                fn(ev[max(k - 2, 0)])
            torch.cuda.synchronize()
            return float(np.mean([x.elapsed_time(y) for x, y in ev]))

        def entry(name, ms, nbytes, note):
            gbs = nbytes / (ms * 1e-3) / 1e9
            out[name] = {"avg_launch_ms": ms, "algorithmic_bytes_per_launch": int(nbytes), "achieved": gbs, "unit": "GB/s",
                         "frac": gbs / peak, "note": note}

        lo = yv3.YOLOv3Loss(rectified_coord_num=self.w["rect"], **self.spec.loss_args)
        lo.auto_all_reduce = False
        ms = timed(lambda e: lo.loss_and_grad(self.td, self.pd, grad_out=self.grad, events=e))
        entry("loss_dense_kernel", ms, n * (4 * A * 5 + 4 * A * L + 4 * m * 5) + 4 * owned * C,
              "loss fwd+bwd alone (the reference's training loop): 4A5 read + 4AL gradient + 4RC + labels per image")
        floor = self.pattern_floor_us()
        if floor is not None:
            out["loss_dense_kernel"]["pattern_floor_us"] = floor
            out["loss_dense_kernel"]["frac_of_pattern_floor"] = floor / (ms * 1e3)
            out["loss_dense_kernel"]["pattern_floor_note"] = (
                "tools/membench.cu on a B200 (profiles/r01_membench.txt): the access pattern alone -- 32-byte reads at a "
                "record stride plus a streaming write of the tensor, no arithmetic; a 20-byte read costs a whole 128-byte "
                "DRAM line and reads and writes do not overlap for this mix")

        def score(e):
            e[0].record()
            self.post.filter(self.pd, self.w["score_thresh"], out=(self.cand, self.counts), sweep_done=e[1])
        ms = timed(score)
        entry("score_tile_kernel", ms, n * 4 * A * L, "decode + score + threshold sweep alone (inference): 4AL read per image")

        dec = yv3.YOLOv3Decoder(self.spec.head_grid_sizes, C, self.spec.anchor_boxes)
        keep_alive = []

        def decode(e):
            keep_alive.clear()
            e[0].record()
            keep_alive.append(dec.decode(self.pd))
            e[1].record()
        try:
            ms = timed(decode)
            entry("decode_kernel", ms, n * (4 * A * L + 4 * A * (L + 8)),
                  "materialising decode (the reference's YOLOv3Decoder.decode contract): 4AL read + 4A(L+8) written; the "
                  "events include the allocation of the nine output tensors by torch's caching allocator")
        except Exception as e:                                     # noqa: BLE001
            out["decode_kernel"] = {"error": type(e).__name__}
        # the un-packed layout (the three head tensors before yolov3_detector.py:80-85 concatenates them): the one-sweep
        # kernel once per head tensor (yv3_loss_filter_fused_heads), a whole step on one stream, against the packed call
        try:
            heads = []
            for h in range(3):
                gh, gw = self.spec.head_grid_sizes[h]
                s0, s1 = self.spec.chan_base[h], self.spec.chan_base[h] + self.spec.head_channels[h]
                heads.append(self.pd[..., s0:s1].reshape(n, gh, gw, -1).contiguous())
            hgrads = [torch.empty_like(x) for x in heads]
            lo2 = yv3.YOLOv3Loss(rectified_coord_num=self.w["rect"], **self.spec.loss_args)
            lo2.auto_all_reduce = False

            def step_heads(e):
                e[0].record()
                lo2.loss_grad_and_filter_heads(self.td, heads, self.post, self.w["score_thresh"], grads_out=hgrads,
                                               filter_out=(self.cand, self.counts))
                self.post.nms(self.cand, self.counts, self.w["nms_thresh"], out=self.keep)
                e[1].record()

            def step_packed(e):
                e[0].record()
                lo.loss_grad_and_filter(self.td, self.pd, self.post, self.w["score_thresh"], grad_out=self.grad,
                                        filter_out=(self.cand, self.counts))
                self.post.nms(self.cand, self.counts, self.w["nms_thresh"], out=self.keep)
                e[1].record()
            ms_h, ms_p = timed(step_heads), timed(step_packed)
            out["heads_layout_step"] = {"ms_per_step_heads": ms_h, "ms_per_step_packed": ms_p, "fused_sweep": bool(lo2.last_fused),
                                        "note": "one whole step (assignment, sweep(s), fold, gather, sort, NMS) issued serially on one "
                                                "stream: three head tensors through yv3_loss_filter_fused_heads vs the packed tensor "
                                                "through yv3_loss_filter_fused; the heads form saves the concatenate copy upstream"}
            del heads, hgrads
        except Exception as e:                                     # noqa: BLE001
            out["heads_layout_step"] = {"error": "%s: %s" % (type(e).__name__, str(e)[:120])}
        return out, owned

    def pattern_floor_us(self):
        """the measured floor of the loss-only access pattern at cfg2 (profiles/r01_membench.txt); other shapes: None"""
        return 61.5 if self.name == "cfg2_416_c80_b64" else None

    # -- N > 1: the sharded run against the single-GPU full batch ------------------------------------------------------------
    def sharding_check(self):
        """Outside the timed region rank 0 regenerates every rank's shard (seeds 1234 + r / 4321 + r), runs the GLOBAL
        batch on one GPU, and compares: the all-reduced 20 loss floats (rtol 1e-6: per-rank partial sums are folded in
        another order) and rank 0's own gradient shard (bit for bit).  yolov3_loss.py:111-112: the mean runs over the
        whole batch."""
        import torch.distributed as dist
        torch, yv3 = self.torch, self.yv3
        n, world = self.n, self.world
        sharded = yv3.YOLOv3Loss(rectified_coord_num=self.w["rect"], **self.spec.loss_args)
        total, grad = sharded.loss_and_grad(self.td, self.pd)        # NCCL all-reduce of the 20 floats inside
        torch.cuda.synchronize()
        res = None
        if self.rank == 0:
            try:
                ps, ts = [], []
                for r in range(world):
                    p_np, t_np = synth(self.spec, n, self.m, r)
                    ps.append(torch.from_numpy(p_np))
                    ts.append(torch.from_numpy(t_np))
                pfull, tfull = torch.cat(ps).to(self.dev), torch.cat(ts).to(self.dev)
                full = yv3.YOLOv3Loss(rectified_coord_num=self.w["rect"], **self.spec.loss_args)
                full.global_batch, full.auto_all_reduce = n * world, False
                tot_f, grad_f = full.loss_and_grad(tfull, pfull)
                a, b = sharded.loss_detail.double(), full.loss_detail.double()
                rel = float(((a - b).abs() / b.abs().clamp_min(1e-30))[b.abs() > 1e-12].max()) if bool((b.abs() > 1e-12).any()) else 0.0
                same = bool(torch.equal(grad, grad_f[:n]))
                res = {"ok": bool(rel <= 1e-6 and same), "max_rel": rel, "grad_shard_bit_exact": same,
                       "global_batch": n * world, "current_num_equal": sharded.current_num == full.current_num}
                del pfull, tfull, grad_f
            except Exception as e:                                 # noqa: BLE001
                res = {"ok": False, "error": "%s: %s" % (type(e).__name__, str(e)[:200])}
        dist.barrier()
        return res


def short_entry(arm, steps, peak):
    """one extra workload: value, ms/step, the dominant kernel's roofline fraction (BASELINE configs[2], [4])"""
    for _ in range(3):
        arm.run(min(steps, arm.unroll))
    arm.fence()
    ms, timed = arm.time_steps(steps, min_ms=30.0)
    sweep_ms = arm.time_sweep(10)
    nbytes, kernel = arm.alg_bytes()
    arm.pipe.check_status()
    return {"workload": arm.name, "desc": arm.w["desc"], "per_gpu_batch": arm.n, "global_batch": arm.n * arm.world,
            "scaling": "strong" if "global_batch" in arm.w else "weak",
            "value": arm.world * arm.n / (ms * 1e-3), "unit": "images/s", "ms_per_step": ms, "timed_steps": timed,
            "fused_sweep": bool(arm.pipe.last_fused) if arm.args.fused else False,
            "roofline": {"kernel": kernel, "avg_launch_ms": sweep_ms, "algorithmic_bytes_per_launch": int(nbytes),
                         "achieved": nbytes / (sweep_ms * 1e-3) / 1e9, "unit": "GB/s", "frac": nbytes / (sweep_ms * 1e-3) / 1e9 / peak},
            "candidates_per_image": float(arm.counts.sum().item()) / arm.n}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg2_416_c80_b64", choices=sorted(WORKLOADS))
    ap.add_argument("--cpu-budget", type=float, default=15.0, help="seconds of CPU work for cpu_baseline")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--fused", type=int, default=1,
                    help="1: FusedPipeline (loss fwd+bwd and the filter on one read of the head tensor, three streams); "
                         "0: loss_and_grad, filter, nms as separate calls on one stream")
    ap.add_argument("--fused-depth", type=int, default=3, help="workspace pairs the pipeline alternates between")
    ap.add_argument("--graph-steps", type=int, default=16, help="steps captured per CUDA graph")
    ap.add_argument("--no-graph", action="store_true", help="time eager launches instead of CUDA-graph replays")
    ap.add_argument("--reduce", default="peer", choices=["peer", "nccl"],
                    help="N > 1: peer = the peer-memory all-reduce kernel inside every step (falls back to nccl when the "
                         "peers cannot be mapped); nccl = one NCCL all-reduce per graph replay, outside the graph")
    ap.add_argument("--min-timed-ms", type=float, default=100.0)
    ap.add_argument("--no-extras", action="store_true", help="skip extra_workloads, roofline_stages, probes")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    w = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return run_reference(args, w, build_spec(w), rank, world)

    import torch
    import torch.distributed as dist
    import yolov3_tensorflow_b200 as yv3

    assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU path in the product)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    reducer, reduce_mode = None, None
    numa = bind_to_gpu_numa_node(local_rank) if world > 1 else None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        if args.reduce == "peer" and args.fused:
            reducer = yv3.PeerAllReduce.create(cap=512)
        reduce_mode = "peer-memory kernel inside every step (yv3_peer_allreduce)" if reducer is not None else \
            "NCCL all-reduce, once per graph replay (outside the graph)"
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"

    arm = Arm(args.workload, args, rank, world, dev, reducer)
    n, m = arm.n, arm.m
    sampler = ClockSampler(local_rank)
    sampler.start()
    sampler.mark = "warmup"
    for _ in range(args.warmup):
        arm.run(1 if args.no_graph else min(arm.unroll, max(1, args.steps)))
    arm.fence()
    sampler.mark = "kernels"
    sweep_ms = arm.time_sweep(min(max(args.steps, 5), 50))
    arm.fence()
    sampler.mark = "timed"
    ms_per_step, timed_steps = arm.time_steps(args.steps, args.min_timed_ms)
    sampler.mark = "after"
    arm.pipe.check_status()
    if reducer is not None:
        reducer.check_status()
    cand_per_image = float(arm.counts.sum().item()) / n
    kept_per_image = float(sum(int(arm.keep[i, :int(arm.counts[i].sum())].sum()) for i in range(min(n, 8)))) / min(n, 8)
    fused_ran = bool(arm.pipe.last_fused) if args.fused else False

    sampler.mark = "e2e"
    e2e_steps = min(args.steps, 100)
    e2e_ms = arm.time_e2e(e2e_steps)
    probe = arm.h2d_probe() if not args.no_extras else None
    sampler.stop_flag = True
    sampler.join(timeout=2)
    clocks = sampler.summary()
    if world > 1:
        # every rank watched its OWN GPU: the line carries the slowest median and the union of the throttle reasons
        allc = [None] * world
        dist.all_gather_object(allc, clocks)
        meds = [c["sm_mhz"] for c in allc if c.get("sm_mhz") is not None]
        clocks = dict(clocks, sm_mhz=min(meds) if meds else None,
                      reasons=sorted({r for c in allc for r in c.get("reasons", [])}),
                      per_rank=[{"sm_mhz": c.get("sm_mhz"), "mem_mhz": c.get("mem_mhz"), "reasons": c.get("reasons")} for c in allc])

    stages, owned, check = None, 0, None
    if not args.no_extras:
        stages, owned = arm.stage_rooflines(peak)
    if world > 1:
        check = arm.sharding_check()
    extras = []
    if not args.no_extras:
        mode = arm.launch_mode()
        del arm.graphs
        arm.graphs = {}
        for name in EXTRA_WORKLOADS:
            if name == args.workload or WORKLOADS[name].get("global_batch", world) % world:
                continue
            torch.cuda.empty_cache()
            try:
                extra = Arm(name, args, rank, world, dev, reducer, pin=False)
                extras.append(short_entry(extra, 32, peak))
                del extra
            except Exception as e:                                 # noqa: BLE001
                extras.append({"workload": name, "error": "%s: %s" % (type(e).__name__, str(e)[:200])})
    else:
        mode = arm.launch_mode()

    if rank == 0:
        nbytes, kernel_name = arm.alg_bytes(owned)
        achieved = nbytes / (sweep_ms * 1e-3) / 1e9
        traffic, traffic_note = None, None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            tj = json.load(open(tp))
            traffic = tj.get(args.workload, {}).get(kernel_name + "_dram_bytes_per_launch")
            sha = git_sha()
            from yolov3_tensorflow_b200 import build as yb
            now_hash = yb.sweep_sources_hash()
            # stale = the sweep kernels' sources changed since the capture (a commit that leaves them alone does not count)
            traffic_note = {"captured_at_sha": tj.get("sha"), "captured_on": tj.get("date"), "bench_sha": sha,
                            "sweep_sources_hash": now_hash, "captured_sources_hash": tj.get("sweep_sources_hash"),
                            "stale": tj.get("sweep_sources_hash") != now_hash,
                            "how": "dram__bytes_read.sum + dram__bytes_write.sum of one ncu --set full capture per workload "
                                   "and kernel (tools/traffic_capture.sh + tools/traffic_json.py regenerate "
                                   "profiles/traffic.json)"}
        A, L = arm.spec.anchors_per_image, arm.spec.box_len
        line = {
            "metric": "head loss fwd+bwd + decode+filter+NMS throughput",
            "value": world * n / (ms_per_step * 1e-3), "unit": "images/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
            "timed_steps": timed_steps, "timed_region_ms": ms_per_step * timed_steps,
            "ms_per_step_per_rank": getattr(arm, "per_rank_ms", None),
            "higher_is_better": True, "scaling": "strong" if "global_batch" in w else "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": config_block(args.workload, w, n, world, {
                "parallelism": "batch-sharded dp%d; the 20 loss floats of every step summed over the ranks: %s" % (world, reduce_mode)
                               if world > 1 else "dp1",
                "l2": "inputs larger than L2: predicts %.0f MB + gradient %.0f MB per GPU vs 126 MB L2" % (
                    arm.pd.numel() * 4 / 1e6, arm.grad.numel() * 4 / 1e6),
                "candidates_per_image": cand_per_image, "kept_per_image": kept_per_image,
                "layout": "packed", "fused_sweep": fused_ran, "launch": mode,
                "schedule": ("FusedPipeline: three streams (sweeps back to back; fold of step k and assignment of step k+depth; "
                             "gather/sort/NMS of step k at high priority), %d workspace pairs" % len(arm.pipe.pairs))
                if args.fused else "separate calls on one stream"}),
            "stage_ms": {kernel_name: sweep_ms, "step": ms_per_step,
                         "note": "the sweep alone from eager launches with events; ms_per_step from the timed region, where "
                                 "the small kernels run under the next step's sweep"},
            "roofline": {"kernel": kernel_name, "bound": "hbm", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_note,
                         "peak_source": peak_src, "algorithmic_bytes_per_launch": int(nbytes), "avg_launch_ms": sweep_ms,
                         "whole_step_frac": nbytes / (ms_per_step * 1e-3) / 1e9 / peak,
                         "dense_2x4AL_bytes_per_launch_for_context": 8 * A * L * n},
            "roofline_stages": stages,
            "e2e": {"value": world * n / (e2e_ms * 1e-3), "unit": "images/s",
                    "h2d_bytes_per_step": arm.e2e_h2d, "d2h_bytes_per_step": arm.e2e_d2h,
                    "returned": "20 loss floats, per-head counts, keep flags, candidate rows and the resized kept rows "
                                "(yv3_collect_kept) of every image; the gradient (%.0f MB) stays on the device for the "
                                "model's backward" % (arm.grad.numel() * 4 / 1e6),
                    "pipelining": "double-buffered: the copy of step k+1 overlaps the compute of step k; results of "
                                  "step k-1 are read on the host every step",
                    "ms_per_step": e2e_ms, "steps": e2e_steps, "h2d_probe": probe, "numa": numa,
                    "frac_of_h2d_probe": (arm.e2e_h2d / (e2e_ms * 1e-3) / 1e9) / probe["per_rank_gbs"] if probe else None},
            # host launches of OUR kernels per step.  Fused: assign, fused sweep, fold, gather, sort (+ the peer all-reduce
            # at N > 1); the two queue-driven NMS kernels are host launches only with <= 8 classes -- with more the sort
            # kernel launches them from the device, and only when their queues hold work (not counted here)
            "gpu_launches": ((5 + (0 if max(arm.spec.class_num, 1) > 8 else 2) + (1 if reducer is not None else 0))
                             if args.fused else (6 + (0 if max(arm.spec.class_num, 1) > 8 else 2))) * timed_steps,
            "gpu_launches_note": "host-side launches of this library's kernels in the timed region; NMS kernels the sort "
                                 "kernel tail-launches from the device are extra",
            "clocks": clocks,
            "extra_workloads": extras,
        }
        if check is not None:
            line["sharding_check"] = check
        if world == 1 and not args.no_cpu_baseline:
            v, s, cores = time_cpu_port(build_spec(w), arm.p_np, arm.t_np, w, args.cpu_budget)   # the oracle enters here only
            line["cpu_baseline"] = {"value": v, "unit": "images/s", "cores": cores, "kind": "port",
                                    "sample": "%d of %d images, one pass (oracle port: torch-CPU fp32 loss + autograd, "
                                              "NumPy decode/filter, Python NMS as run.py runs it)" % (s, n)}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
