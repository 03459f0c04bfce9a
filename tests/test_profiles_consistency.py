"""The committed evidence must describe the committed code: profiles/traffic.json carries a content hash of the sweep
kernels' sources (bench.py marks its DRAM-traffic figure stale when they differ) -- a change to those sources without a
new capture (tools/traffic_capture.sh + tools/traffic_json.py) fails here, on the CPU, before a stale number is published."""
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_traffic_json_matches_the_sweep_sources():
    from yolov3_tensorflow_b200 import build
    with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
        tj = json.load(f)
    assert tj.get("sweep_sources_hash") == build.sweep_sources_hash(), \
        "the sweep kernels changed since profiles/traffic.json was captured: run tools/traffic_capture.sh on a B200"
    for workload in ("cfg2_416_c80_b64", "cfg3_608_c80_b64", "cfg5_512_c0_g128"):
        assert tj[workload]["fused_sweep_kernel_dram_bytes_per_launch"] > 0
    assert tj["cfg2_416_c80_b64"]["loss_dense_kernel_dram_bytes_per_launch"] > 0
    assert tj["cfg2_416_c80_b64"]["decode_kernel_dram_bytes_per_launch"] > 0


def test_committed_bench_lines_carry_the_contract_keys():
    need = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
            "dtype", "data", "config", "roofline", "e2e", "gpu_launches", "clocks"}
    for name in ("r02_bench_n1.json", "r02_bench_n2.json", "r02_bench_n4.json", "r02_bench_n8.json"):
        with open(os.path.join(ROOT, "profiles", name)) as f:
            line = json.load(f)
        assert need <= set(line), (name, sorted(need - set(line)))
        assert line["config"]["workload"] == "cfg2_416_c80_b64"
        assert set(line["roofline"]) >= {"bound", "achieved", "peak", "unit", "frac", "traffic"}
        assert set(line["e2e"]) >= {"value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"}
        assert not set(line["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
        if line["n_gpus"] > 1:
            assert line["sharding_check"]["ok"] and line["sharding_check"]["grad_shard_bit_exact"]
    with open(os.path.join(ROOT, "profiles", "r02_bench_n1.json")) as f:
        assert "cpu_baseline" in json.load(f)
