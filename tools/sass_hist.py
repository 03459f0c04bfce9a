"""Developer helper: executed-instruction / stall-sample histogram over blocks of SASS lines (ncu --page source --csv)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
B = int(sys.argv[2]) if len(sys.argv) > 2 else 60
hdr = rows[1]; data = rows[2:]
iS = hdr.index('# Samples'); iSrc = hdr.index('Source'); iE = hdr.index('Instructions Executed')
tot = sum(int(r[iS]) for r in data); E = sum(int(r[iE]) for r in data)
print('samples', tot, 'executed', E)
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
agg = {s: sum(int(r[hdr.index(s)]) for r in data) for s in stalls}
print(sorted(agg.items(), key=lambda x: -x[1])[:8])
for b in range(0, len(data), B):
    blk = data[b:b + B]; e = sum(int(r[iE]) for r in blk); s = sum(int(r[iS]) for r in blk)
    if e > 0.01 * E or s > 0.01 * tot:
        ops = {}
        for r in blk:
            t = r[iSrc].split()
            op = t[1] if t[0].startswith('@') else t[0]
            op = op.split('.')[0]; ops[op] = ops.get(op, 0) + int(r[iE])
        top = sorted(ops.items(), key=lambda x: -x[1])[:7]
        print('%5d-%5d exec=%8d (%4.1f%%) samples=%5d (%4.1f%%) %s' % (b, b + B - 1, e, 100 * e / E, s, 100 * s / tot,
              ' '.join('%s:%d' % (k, v // 1000) for k, v in top)))
if len(sys.argv) > 3:
    idx = sorted(range(len(data)), key=lambda i: -int(data[i][iS]))[:int(sys.argv[3])]
    for i in sorted(idx):
        r = data[i]
        top = sorted(((int(r[hdr.index(s)]), s) for s in stalls), reverse=True)[:2]
        print(i, r[iS], r[iE], r[iSrc][:64], top)
