#!/usr/bin/env python
"""Record the measured DRAM traffic of one chain-kernel launch for bench.py's `roofline.traffic`:

    python tools/ncu_traffic.py WORKLOAD REPORT.ncu-rep SUMMARY_NAME [SRC_SHA]

Reads dram__bytes_read.sum / dram__bytes_write.sum of the first spa_chain_kernel in the `ncu --set full` report and
stores them in profiles/ncu_traffic.json together with the hash of the kernel sources the capture belongs to
(bench.kernel_source_sha(); pass SRC_SHA when the capture was taken from another tree state).  bench.py reports the
figure only while that hash matches the sources it is timing."""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    wl, rep, summary = sys.argv[1:4]
    if len(sys.argv) > 4:
        sha = sys.argv[4]
    else:
        import bench
        sha = bench.kernel_source_sha()
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    for vals in rows[2:]:
        if "spa_chain_kernel" not in vals[hdr.index("Kernel Name")]:
            continue
        get = lambda m: float(vals[hdr.index(m)].replace(",", "")) * UNIT[units[hdr.index(m)]]
        rec = {"dram_read": int(get("dram__bytes_read.sum")), "dram_write": int(get("dram__bytes_write.sum")),
               "gpu_time_us_under_ncu": get_time(vals, hdr, units), "src_sha": sha, "summary": summary}
        path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
        try:
            t = json.load(open(path))
        except (OSError, ValueError):
            t = {"workloads": {}}
        t["workloads"][wl] = rec
        json.dump(t, open(path, "w"), indent=1, sort_keys=True)
        print(wl, rec)
        return
    raise SystemExit("no spa_chain_kernel launch in the report")


def get_time(vals, hdr, units):
    i = hdr.index("gpu__time_duration.sum")
    scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}.get(units[i], 1.0)
    return float(vals[i].replace(",", "")) * scale


if __name__ == "__main__":
    main()
