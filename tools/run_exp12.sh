w=cfg5_512_c0_b16
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_$w.csv python tools/profile_step.py $w 2 > /dev/null 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open('gpurun_out/launches_$w.csv')) if len(r)>5]
hdr=rows[0]; i=hdr.index('Kernel Name'); j=hdr.index('Metric Value'); g=hdr.index('Grid Size'); b=hdr.index('Block Size')
for r in rows[-9:]: print('  ', r[i][:50], r[g], r[b], r[j])
PY
