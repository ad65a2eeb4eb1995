#!/usr/bin/env python3
"""Record the DRAM traffic of a workload's dominant kernel from an `ncu --set full` report into
profiles/traffic.json (bench.py's roofline.traffic reads it; nothing is hard-coded in bench.py).
usage: tools/ncu_traffic.py <workload> <report.ncu-rep> <kernel-name-regex> [commit]"""
import csv
import io
import json
import os
import subprocess
import sys

wl, rep, rx = sys.argv[1], sys.argv[2], sys.argv[3]
commit = sys.argv[4] if len(sys.argv) > 4 else subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
raw = subprocess.run(["ncu", "-i", rep, "--kernel-name", "regex:" + rx, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
best = None
for vals in rows[2:]:
    g = lambda k: vals[hdr.index(k)]  # noqa: E731
    u = lambda k: units[hdr.index(k)]  # noqa: E731

    def to_bytes(k):
        v, un = float(g(k).replace(",", "")), u(k).lower()
        return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(un, 1)

    def to_us(k):
        v, un = float(g(k).replace(",", "")), u(k).lower()
        return v * {"ns": 1e-3, "us": 1, "usecond": 1, "ms": 1e3, "msecond": 1e3, "nsecond": 1e-3, "second": 1e6}.get(un, 1)
    rec = {"kernel": g("Kernel Name"), "dram_read_bytes": to_bytes("dram__bytes_read.sum"),
           "dram_write_bytes": to_bytes("dram__bytes_write.sum"), "capture_duration_us": to_us("gpu__time_duration.sum"),
           "file": os.path.basename(rep), "commit": commit}
    if best is None or rec["capture_duration_us"] > best["capture_duration_us"]:
        best = rec
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
path = os.environ.get("TRAFFIC_JSON", os.path.join(root, "profiles", "traffic.json"))
try:
    db = json.load(open(path))
except (OSError, ValueError):
    db = {}
db[wl] = best
json.dump(db, open(path, "w"), indent=1, sort_keys=True)
print(wl, best)
