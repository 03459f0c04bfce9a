"""Worker of tests/test_gpu_multi_rank.py (launched by torchrun, one process per GPU, NCCL): the sharded product path
against the single-GPU full-batch call (yolov3_loss.py:111-112: the batch mean runs over the WHOLE batch)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)

from helpers import make_predicts, make_spec, make_targets, product_loss_args  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=dev)
    import yolov3_tensorflow_b200 as yv3

    for name, per_rank, rect in (("coco_64x96_c4", 3, True), ("coco_416_c80", 2, False), ("single_160_c1_focal", 4, True)):
        spec = make_spec(name)
        yv3.backend.set_epsilon(spec.eps)
        n_full, m = per_rank * world, 6
        p = make_predicts(spec, n_full, seed=77)                     # every rank builds the same global batch
        t = make_targets(spec, n_full, m, seed=78, min_valid=0, pad_in_middle=True)
        lo, hi = rank * per_rank, (rank + 1) * per_rank
        rcn = n_full if rect else -1                                  # two active steps, then off
        sharded = yv3.YOLOv3Loss(rectified_coord_num=rcn, **product_loss_args(spec))
        full = yv3.YOLOv3Loss(rectified_coord_num=rcn, **product_loss_args(spec))
        full.global_batch, full.auto_all_reduce = n_full, False      # the single-GPU call on the whole batch
        pd, td = torch.from_numpy(p).to(dev), torch.from_numpy(t).to(dev)
        for step in range(3):
            tot_s, grad_s = sharded.loss_and_grad(td[lo:hi], pd[lo:hi])
            tot_f, grad_f = full.loss_and_grad(td, pd)
            sharded.check_status()
            # this rank's gradient shard: the same per-record arithmetic with the same 1 / N_global -> the same bits
            assert torch.equal(grad_s, grad_f[lo:hi]), "%s step %d: gradient shard differs" % (name, step)
            # the all-reduced detail: per-rank partial sums folded in another order (fp32 noise only)
            np.testing.assert_allclose(sharded.loss_detail.cpu().numpy(), full.loss_detail.cpu().numpy(), rtol=2e-6, atol=1e-7)
            np.testing.assert_allclose(tot_s.item(), tot_f.item(), rtol=2e-6)
            assert sharded.current_num == full.current_num, (sharded.current_num, full.current_num)
        assert sharded.current_num == (2 * n_full if rect else 0)

    # ragged shards: rank r holds r + 2 images.  Without global_batch every rank raises (no silent wrong mean) ...
    spec = make_spec("coco_64x96_c4")
    sizes = [r + 2 for r in range(world)]
    n_full = sum(sizes)
    p = make_predicts(spec, n_full, seed=91)
    t = make_targets(spec, n_full, 6, seed=92, min_valid=1)
    lo = sum(sizes[:rank])
    hi = lo + sizes[rank]
    pd, td = torch.from_numpy(p).to(dev), torch.from_numpy(t).to(dev)
    ragged = yv3.YOLOv3Loss(rectified_coord_num=-1, **product_loss_args(spec))
    try:
        ragged.loss_and_grad(td[lo:hi], pd[lo:hi])
        raise AssertionError("ragged shards were accepted without global_batch")
    except ValueError:
        pass
    # ... and with it the result is the full batch's
    ragged.global_batch = n_full
    tot_s, grad_s = ragged.loss_and_grad(td[lo:hi], pd[lo:hi])
    full = yv3.YOLOv3Loss(rectified_coord_num=-1, **product_loss_args(spec))
    full.global_batch, full.auto_all_reduce = n_full, False
    tot_f, grad_f = full.loss_and_grad(td, pd)
    assert torch.equal(grad_s, grad_f[lo:hi])
    np.testing.assert_allclose(ragged.loss_detail.cpu().numpy(), full.loss_detail.cpu().numpy(), rtol=2e-6, atol=1e-7)

    # the pipeline's rows, all-reduced once for all steps
    spec = make_spec("coco_96x64_c13")
    per_rank, k = 3, 6
    n_full = per_rank * world
    ps = [make_predicts(spec, n_full, seed=200 + i) for i in range(k)]
    ts = [make_targets(spec, n_full, 5, seed=300 + i, min_valid=0) for i in range(k)]
    lo, hi = rank * per_rank, (rank + 1) * per_rank
    pipe = yv3.FusedPipeline(product_loss_args(spec), 0.05, 0.45, depth=3, rectified_coord_num=2 * n_full)
    shard_p = [torch.from_numpy(x[lo:hi]).to(dev) for x in ps]
    shard_t = [torch.from_numpy(x[lo:hi]).to(dev) for x in ts]
    rows = torch.zeros((k, 20), device=dev)
    grads = [torch.empty_like(x) for x in shard_p]
    cand = torch.zeros((per_rank, pipe.capacity, 8), device=dev)
    counts = torch.zeros((per_rank, 3), dtype=torch.int32, device=dev)
    keep = torch.zeros((per_rank, pipe.capacity), dtype=torch.uint8, device=dev)
    pipe.run(shard_t, shard_p, grads, rows, cand, counts, keep)
    pipe.all_reduce_rows(rows)
    full = yv3.YOLOv3Loss(rectified_coord_num=2 * n_full, **product_loss_args(spec))
    full.global_batch, full.auto_all_reduce = n_full, False
    for i in range(k):
        tot_f, grad_f = full.loss_and_grad(torch.from_numpy(ts[i]).to(dev), torch.from_numpy(ps[i]).to(dev))
        assert torch.equal(grads[i], grad_f[lo:hi]), "pipeline step %d: gradient shard differs" % i
        np.testing.assert_allclose(rows[i, :18].cpu().numpy(), full.loss_detail.reshape(-1).cpu().numpy(), rtol=2e-6, atol=1e-7)
    assert pipe.current_num == full.current_num == 3 * n_full
    # ---- the peer-memory all-reduce kernel (csrc/peer_reduce.cu) against NCCL, then inside the pipeline and a CUDA graph
    red = yv3.PeerAllReduce.create(cap=512)
    peer = "nccl only (peers could not be mapped)"
    if red is not None:
        peer = "peer kernel"
        g = torch.Generator(device=dev).manual_seed(1000 + rank)
        for it in range(40):                                           # both parities, many epochs, varying sizes
            cnt = [20, 320, 1, 512][it % 4]
            x = torch.randn(cnt, generator=g, device=dev)
            want = x.clone()
            dist.all_reduce(want)
            got = red(x.clone())
            np.testing.assert_allclose(got.cpu().numpy(), want.cpu().numpy(), rtol=1e-6, atol=1e-6)
            gathered = [torch.empty_like(got) for _ in range(world)]
            dist.all_gather(gathered, got)
            assert all(torch.equal(gathered[0], t) for t in gathered), "ranks hold different sums"
        red.check_status()
        pipe2 = yv3.FusedPipeline(product_loss_args(spec), 0.05, 0.45, depth=3, rectified_coord_num=-1, reducer=red)
        rows2 = torch.zeros((k, 20), device=dev)
        graph = pipe2.capture(shard_t, shard_p, grads, rows2, cand, counts, keep)
        full2 = yv3.YOLOv3Loss(rectified_coord_num=-1, **product_loss_args(spec))
        full2.global_batch, full2.auto_all_reduce = n_full, False
        for rep in range(3):
            rows2.zero_()
            graph.replay()
            torch.cuda.synchronize()
            for i in range(k):
                full2.loss_and_grad(torch.from_numpy(ts[i]).to(dev), torch.from_numpy(ps[i]).to(dev))
                np.testing.assert_allclose(rows2[i, :18].cpu().numpy(), full2.loss_detail.reshape(-1).cpu().numpy(),
                                           rtol=2e-6, atol=1e-7)
        red.check_status()
    dist.barrier()
    if rank == 0:
        print("multi-rank worker ok: world %d, %s" % (world, peer))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
