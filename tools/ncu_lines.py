"""Developer helper: instructions executed and stall samples per SOURCE LINE from an .ncu-rep captured with
--import-source on (built with -lineinfo).   python tools/ncu_lines.py prof.ncu-rep [top]"""
import collections
import csv
import subprocess
import sys


def main(path, top=40):
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    cur, hdr = None, None
    agg, smp, src = collections.Counter(), collections.Counter(), {}
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur = r[1].split("/")[-1]
        elif r[0] == "Line No":
            hdr = r
        elif cur and hdr and r[0].isdigit() and len(r) > 8:
            k = (cur, int(r[0]))
            try:
                agg[k] += int(r[hdr.index("Instructions Executed")] or 0)
                smp[k] += int(r[hdr.index("# Samples")] or 0)
            except ValueError:
                pass
            src[k] = r[1]
    tot, stot = sum(agg.values()), max(1, sum(smp.values()))
    byfile = collections.Counter()
    for k, v in agg.items():
        byfile[k[0]] += v
    print("warp instructions:", tot, dict(byfile))
    for k, v in agg.most_common(int(top)):
        print("%6.2f%% inst %6.2f%% samples  %s:%d  %s" % (100.0 * v / tot, 100.0 * smp[k] / stot, k[0], k[1], src[k].strip()[:100]))


if __name__ == "__main__":
    main(*sys.argv[1:])
