"""Developer probe: which of the small kernels costs the sweep how much.  The fused schedule of FusedPipeline with parts
switched off (results are then wrong on purpose): graph of 16 steps, replayed; prints us/step per variant."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import yolov3_tensorflow_b200 as yv3  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "cfg2_416_c80_b64"
    w = bench.WORKLOADS[name]
    spec = bench.product_spec(w)
    yv3.backend.set_epsilon(spec.eps)
    dev = torch.device("cuda", 0)
    n, m = bench.per_rank_batch(w, 1), w["m"]
    p_np, t_np = bench.synth(spec, n, m, 0)
    pd, td = torch.from_numpy(p_np).to(dev), torch.from_numpy(t_np).to(dev)
    pipe = yv3.FusedPipeline(spec.loss_args, w["score_thresh"], w["nms_thresh"], depth=3, rectified_coord_num=w["rect"])
    grad = torch.empty_like(pd)
    cap = pipe.capacity
    cand = torch.empty((n, cap, 8), device=dev)
    counts = torch.empty((n, 3), dtype=torch.int32, device=dev)
    keep = torch.empty((n, cap), dtype=torch.uint8, device=dev)
    rows = torch.zeros((16, 20), device=dev)
    pipe.run(td, pd, grad, rows, cand, counts, keep, count=16)
    torch.cuda.synchronize()
    thr = w["score_thresh"]

    def variant(assign, fold, gather, nms, streams=True):
        cur = torch.cuda.current_stream()
        aux, post = (pipe.aux_stream, pipe.post_stream) if streams else (cur, cur)
        depth = 3

        def part(k, ph):
            lo, po = pipe.pairs[k % depth]
            lo.loss_grad_and_filter(td, pd, po, thr, grad_out=grad, out=rows[k], filter_out=(cand, counts), phases=ph, rect_override=w["rect"] >= 0)
        if streams:
            aux.wait_stream(cur)
            post.wait_stream(cur)
        for k in range(16):
            swept = torch.cuda.Event()
            if assign:
                with torch.cuda.stream(aux):
                    part(k, 1)
                    e = torch.cuda.Event()
                    e.record(aux)
                cur.wait_event(e)
            part(k, 2)
            swept.record(cur)
            with torch.cuda.stream(post):
                post.wait_event(swept)
                if gather:
                    part(k, 8)
                if nms:
                    pipe.pairs[k % depth][1].nms(cand, counts, w["nms_thresh"], out=keep)
            with torch.cuda.stream(aux):
                aux.wait_event(swept)
                if fold:
                    part(k, 4)
        if streams:
            cur.wait_stream(post)
            cur.wait_stream(aux)

    def timeit(fn, label):
        fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn()
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(40):
            g.replay()
        b.record()
        torch.cuda.synchronize()
        print("%-44s %.1f us/step" % (label, 1e3 * a.elapsed_time(b) / (40 * 16)), flush=True)

    timeit(lambda: variant(0, 0, 0, 0), "sweeps only")
    timeit(lambda: variant(1, 0, 0, 0), "+ assign (serialised before its sweep)")
    timeit(lambda: variant(0, 1, 0, 0), "+ fold")
    timeit(lambda: variant(0, 0, 1, 0), "+ gather")
    timeit(lambda: variant(0, 0, 1, 1), "+ gather + sort/nms")
    timeit(lambda: variant(0, 1, 1, 1), "+ fold + gather + sort/nms")
    timeit(lambda: pipe.run(td, pd, grad, rows, cand, counts, keep, count=16), "the pipeline (all)")


if __name__ == "__main__":
    main()
