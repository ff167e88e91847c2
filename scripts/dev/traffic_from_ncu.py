#!/usr/bin/env python
"""profiles/traffic.json entry from an `ncu --set full` report: DRAM bytes read + written by ONE launch of the kernel,
stamped with the digest of the kernel sources it was captured from (bench.py reports `roofline.traffic` only when the
digest still matches).

    python scripts/dev/traffic_from_ncu.py <report.ncu-rep> <workload> <variant>      e.g.  ... prof.ncu-rep c3 tb2q:128x14
"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (kernel_digest)

UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def main():
    rep, workload, variant = sys.argv[1:4]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2]
    total = 0.0
    for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        i = hdr.index(m)
        total += float(vals[i]) * UNIT[units[i]]
    path = os.path.join(ROOT, "profiles", "traffic.json")
    t = json.load(open(path))
    t.setdefault(workload, {})[variant] = {"bytes": int(total), "digest": bench.kernel_digest(), "report": os.path.basename(rep)}
    json.dump(t, open(path, "w"), indent=1)
    print(workload, variant, int(total), bench.kernel_digest())


if __name__ == "__main__":
    main()
