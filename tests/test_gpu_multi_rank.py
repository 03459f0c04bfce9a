"""GPU, world size >= 2 over NCCL: the sharded product path (YOLOv3Loss with torch.distributed, FusedPipeline rows)
equals the single-GPU full-batch call -- gradient shards bit for bit, the all-reduced 20 floats to fp32 fold noise, the
rectified counter, ragged shards refused unless global_batch is given.  Skipped on a one-GPU box (bench.py --gpus N prints
the same comparison as "sharding_check" so the driver's scaling run carries it)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu


def test_sharded_path_equals_full_batch():
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    here = os.path.dirname(os.path.abspath(__file__))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=%d" % world,
           "--master-addr", "127.0.0.1", "--master-port", "29431", os.path.join(here, "multi_rank_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "multi-rank worker ok" in r.stdout
    print([l for l in r.stdout.splitlines() if "multi-rank worker ok" in l][-1])     # says whether the peer kernel ran
