"""CPU, world_size 2, gloo: the batch-sharding contract of the loss (DESIGN.md "multi-GPU").

Each rank holds a contiguous shard, reduces its share  sum_local(per-image terms) / N_global  and scales
its gradient shard by 1 / N_global; the only collective is the sum all-reduce of the 20 loss floats.
The product kernels need a GPU, so the per-rank share here comes from the oracle; what is tested is that
shares + all-reduce reproduce the single-process batch mean and gradient exactly as the design claims."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from helpers import make_predicts, make_spec, make_targets
from oracle import ref_torch as rt


def _worker(rank, world, port, name, n_global, m, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    spec = make_spec(name)
    p = make_predicts(spec, n_global, seed=7)
    t = make_targets(spec, n_global, m, seed=8, min_valid=1)
    per = n_global // world
    sl = slice(rank * per, (rank + 1) * per)
    # local mean over the shard, then rescale to a share of the global mean (what the kernel's n_global does)
    total, detail, grad = rt.loss_and_grad(t[sl], p[sl], spec, rectified_active=True, dtype=torch.float64)
    share = torch.zeros(20, dtype=torch.float64)
    share[:18] = torch.from_numpy(detail.reshape(-1)) * (per / n_global)
    share[18] = share[:18].sum()
    dist.all_reduce(share, op=dist.ReduceOp.SUM)
    grad_shard = grad * (per / n_global)
    np.save(os.path.join(out_dir, "share%d.npy" % rank), share.numpy())
    np.save(os.path.join(out_dir, "grad%d.npy" % rank), grad_shard)
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_sharded_loss_equals_full_batch(tmp_path):
    name, n_global, m, world = "coco_64x96_c4", 4, 6, 2
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(world, port, name, n_global, m, str(tmp_path)), nprocs=world, join=True)
    spec = make_spec(name)
    p = make_predicts(spec, n_global, seed=7)
    t = make_targets(spec, n_global, m, seed=8, min_valid=1)
    total, detail, grad = rt.loss_and_grad(t, p, spec, rectified_active=True, dtype=torch.float64)
    s0 = np.load(tmp_path / "share0.npy")
    s1 = np.load(tmp_path / "share1.npy")
    assert np.array_equal(s0, s1)                                    # every rank holds the global values
    np.testing.assert_allclose(s0[:18].reshape(6, 3), detail, rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(s0[18], float(total), rtol=1e-12)
    g = np.concatenate([np.load(tmp_path / "grad0.npy"), np.load(tmp_path / "grad1.npy")], 0)
    np.testing.assert_allclose(g, grad, rtol=1e-12, atol=1e-15)
