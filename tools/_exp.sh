timeout 600 python -m pytest tests/test_gpu_nms_edges.py tests/test_gpu_parity.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -3
for s in 0 12; do YV3_NMS_SKIP=$s timeout 120 python tools/nms_probe.py cfg5_nms_stress_b16; done
YV3_NMS_SKIP=8 timeout 120 python tools/nms_probe.py cfg4_416_c80_b256_dense
timeout 300 python bench.py --workload cfg5_nms_stress_b16 --no-cpu-baseline --no-extras 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); print(d['config']['workload'], d['ms_per_step'], d['stage_ms']['fused_sweep_kernel'])"
