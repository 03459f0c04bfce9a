for w in cfg4_416_c80_b256_dense cfg5_512_c0_b16; do
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_$w.csv python tools/profile_step.py $w 2 > /dev/null 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open('gpurun_out/launches_$w.csv')) if len(r)>5]
hdr=rows[0]; i=hdr.index('Kernel Name'); j=hdr.index('Metric Value')
print('$w')
for r in rows[-10:]: print('  ', r[i][:60], r[j])
PY
done
