for v in "256" "128"; do
  export YV3_EXTRA_NVCC="-DYV3_DENSE_THREADS=$v"
  timeout 600 python -c "import __graft_entry__ as g; g.build()" 2>&1 | tail -1
  echo "threads $v"; YV3_CPW=1 timeout 200 python bench.py --no-cpu-baseline --steps 200 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['config']['launch'], d['stage_ms'], d['roofline']['frac'])"
  YV3_CPW=1 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_t$v.csv python tools/profile_step.py > /dev/null 2>&1
  python - <<PY
import csv
rows=[r for r in csv.reader(open('gpurun_out/launches_t$v.csv')) if len(r)>5]
hdr=rows[0]; i=hdr.index('Kernel Name'); j=hdr.index('Metric Value')
for r in rows[-8:]: print(r[i][:50], r[j])
PY
done
