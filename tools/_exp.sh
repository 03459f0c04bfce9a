timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
for s in 0 12; do YV3_NMS_SKIP=$s timeout 120 python tools/nms_probe.py cfg5_nms_stress_b16; done
for w in cfg5_nms_stress_b16 cfg5_512_c0_g128 cfg2_416_c80_b64; do timeout 300 python bench.py --workload $w --no-cpu-baseline --no-extras 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); print(d['config']['workload'], d['ms_per_step'], d['stage_ms']['fused_sweep_kernel'])"; done
for r in 0 4 5 6 7; do YV3_BENCH_SEED_RANK=$r timeout 200 python bench.py --workload cfg3_608_c80_b64 --no-cpu-baseline --no-extras --steps 100 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); print('seed rank $r', round(d['ms_per_step'],4), d['config']['candidates_per_image'], d['config']['kept_per_image'])"; done
