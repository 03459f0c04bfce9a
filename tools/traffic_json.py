"""profiles/traffic.json from the reports tools/traffic_capture.sh brought back (run where ncu is installed, no GPU):
    python tools/traffic_json.py [gpurun_out]
Per workload and kernel: dram__bytes_read.sum + dram__bytes_write.sum of ONE launch (bench.py's roofline.traffic), the
read / write split, the launch's duration under ncu (cold cache, serialised: not a bench value); top level: the commit
the library was built from, the date and the content hash of the sweep kernels' sources, so that bench.py can mark the
figure stale when those kernels changed since."""
import csv
import datetime
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REPORTS = {   # report stem -> (workload, kernel)
    "traffic_cfg2_fused": ("cfg2_416_c80_b64", "fused_sweep_kernel"),
    "traffic_cfg2_dense": ("cfg2_416_c80_b64", "loss_dense_kernel"),
    "traffic_cfg2_decode": ("cfg2_416_c80_b64", "decode_kernel"),
    "traffic_cfg3_fused": ("cfg3_608_c80_b64", "fused_sweep_kernel"),
    "traffic_cfg5_fused": ("cfg5_512_c0_g128", "fused_sweep_kernel"),
}
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "usecond": 1.0,
        "nsecond": 1e-3, "msecond": 1e3, "second": 1e6}


def metrics(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2]

    def get(name):
        i = hdr.index(name)
        return float(vals[i].replace(",", "")) * UNIT[units[i]]
    return {"kernel": vals[hdr.index("Kernel Name")].split("(")[0], "read": get("dram__bytes_read.sum"),
            "write": get("dram__bytes_write.sum"), "us": get("gpu__time_duration.sum")}


def main():
    src = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out")
    sha = subprocess.run(["git", "rev-parse", "--short", "HEAD"], cwd=ROOT, capture_output=True, text=True).stdout.strip()
    sys.path.insert(0, ROOT)
    from yolov3_tensorflow_b200 import build as yb
    tj = {"sha": sha, "date": datetime.date.today().isoformat(), "sweep_sources_hash": yb.sweep_sources_hash(),
          "how": "tools/traffic_capture.sh on one B200: ncu --set full --clock-control none, the launch after three warm-up "
                 "launches of tools/sweep_probe.py --once; dram__bytes_read.sum + dram__bytes_write.sum of that launch"}
    for stem, (workload, kernel) in REPORTS.items():
        path = os.path.join(src, stem + ".ncu-rep")
        if not os.path.exists(path):
            print("missing", path)
            continue
        m = metrics(path)
        assert kernel in m["kernel"], (kernel, m["kernel"])
        w = tj.setdefault(workload, {})
        w[kernel + "_dram_bytes_per_launch"] = int(round(m["read"] + m["write"]))
        w[kernel + "_dram_read_bytes"] = int(round(m["read"]))
        w[kernel + "_dram_write_bytes"] = int(round(m["write"]))
        w[kernel + "_ncu_us"] = round(m["us"], 2)
        print("%-22s %-20s read %8.2f MB  write %8.2f MB  %8.2f us under ncu" % (workload, kernel, m["read"] / 1e6,
                                                                                m["write"] / 1e6, m["us"]))
    with open(os.path.join(ROOT, "profiles", "traffic.json"), "w") as f:
        json.dump(tj, f, indent=1)
        f.write("\n")


if __name__ == "__main__":
    main()
