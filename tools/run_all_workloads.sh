# developer helper: one short bench line per workload (GPU box)
for w in cfg2_416_c80_b64 cfg3_608_c80_b64 cfg4_416_c80_b256_dense cfg5_512_c0_g128 cfg5_nms_stress_b16; do
  timeout 300 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import sys, json
l = sys.stdin.readline()
try:
    d = json.loads(l)
    print(d['config']['workload'], 'img/s %.0f' % d['value'], 'ms/step %.3f' % d['ms_per_step'], d['stage_ms'], 'frac %.3f' % d['roofline']['frac'], 'e2e %.0f' % d['e2e']['value'], d['config']['launch'], 'cand/img %.0f kept %.0f' % (d['config']['candidates_per_image'], d['config']['kept_per_image']))
except Exception as e:
    print('FAILED', l[:300])
"
done
