"""Developer helper: per kernel of lib/libyv3head.so, the SASS mnemonics that show how data moves (cuobjdump -sass):
UBLKCP (1-D bulk / TMA copies), SYNCS (mbarrier), LDG/STG widths, MUFU, DADD/DMUL (fp64 NMS arithmetic) and -- by design
absent -- UTC*MMA / HMMA (no stage of the path is a contraction).   python tools/sass_kernels.py > profiles/r02_sass_hist.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "yolov3-tensorflow_b200", "lib", "libyv3head.so")
WATCH = ["UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "LDG.E.128", "STG.E.128", "LDS.128", "STS.128", "LDGSTS", "MUFU", "DADD", "DMUL", "DFMA",
         "SHFL", "VOTE", "ATOMS", "ATOMG", "RED", "BAR", "UCGABAR", "UTCHMMA", "UTCQMMA", "HMMA", "IMMA", "CALL"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    print("cuobjdump -sass yolov3-tensorflow_b200/lib/libyv3head.so  (arch lines: %s)" %
          sorted(set(re.findall(r"arch = (sm_\w+)", out))))
    name, counts, total = None, None, 0
    rows = []
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            if name:
                rows.append((name, total, counts))
            name, counts, total = m.group(1), collections.Counter(), 0
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and name:
            total += 1
            op = m.group(1)
            for w in WATCH:
                if op.startswith(w):
                    counts[w] += 1
    if name:
        rows.append((name, total, counts))
    dem = subprocess.run(["cu++filt"] + [r[0] for r in rows], capture_output=True, text=True).stdout.splitlines()
    for (nm, tot, c), d in zip(rows, dem):
        short = re.sub(r"\(.*", "", d) if d else nm
        if "yv3::" not in short:
            continue                                               # cudadevrt's own device functions, libdevice slow paths
        print("%-58s %6d instr  %s" % (short[:58], tot, " ".join("%s:%d" % (k, c[k]) for k in WATCH if c[k])))
    allc = collections.Counter()
    for _, _, c in rows:
        allc.update(c)
    print("\nlibrary totals:", " ".join("%s:%d" % (k, allc[k]) for k in WATCH))
    print("tensor-core mnemonics (UTC*MMA / HMMA / IMMA): %d -- none by design: no stage of the head path is a dense contraction"
          % (allc["UTCHMMA"] + allc["UTCQMMA"] + allc["HMMA"] + allc["IMMA"]))


if __name__ == "__main__":
    main()
