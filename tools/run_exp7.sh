timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for d in 0 64; do echo "debug $d"; YV3_DEBUG=$d timeout 120 python tools/dense_bench.py cfg2_416_c80_b64 100 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['loss_dense'], d['loss_all']['median_us'], d['loss_fwd_only']['median_us'])"; done
timeout 200 python bench.py --no-cpu-baseline --steps 200 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['config']['launch'], d['stage_ms'], d['roofline']['frac'])"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_v14.csv python tools/profile_step.py > /dev/null 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open('gpurun_out/launches_v14.csv')) if len(r)>5]
hdr=rows[0]; i=hdr.index('Kernel Name'); j=hdr.index('Metric Value')
for r in rows[-10:]: print(r[i][:50], r[j])
PY
