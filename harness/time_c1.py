#!/usr/bin/env python3
"""CPU-A / drop-in timing of BASELINE config C1 through the harness's DuckDB shell (harness/c1_query.sql).
usage: time_c1.py --flavour cpu|b200 [--threads N] [--out file.json]"""
import argparse
import json
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFDIR = os.path.join(ROOT, "oracle", "_ref")
ap = argparse.ArgumentParser()
ap.add_argument("--flavour", default="cpu", choices=["cpu", "b200"])
ap.add_argument("--threads", type=int, default=len(os.sched_getaffinity(0)))
ap.add_argument("--out", default=None)
a = ap.parse_args()
slices = []
for rep in range(2):
    for lo in range(0, 10000, 600):
        slices.append("INSERT INTO r1 SELECT group_id, (m).r_squared, (m).n_observations, (m).coefficients[1] FROM (SELECT group_id, "
                      "ols_fit_agg(y, [x1, x2, x3]) AS m FROM t WHERE group_id >= %d AND group_id < %d GROUP BY group_id);" % (lo, lo + 600))
    if rep == 0:
        slices.append("DELETE FROM r1;")
n_plain = len(slices)
for lo in range(0, 10000, 600):
    slices.append("INSERT INTO r3 SELECT group_id, (m).r_squared, (m).p_values[1] FROM (SELECT group_id, ols_fit_agg(y, [x1, x2, x3], "
                  "{'compute_inference': true}) AS m FROM t WHERE group_id >= %d AND group_id < %d GROUP BY group_id);" % (lo, lo + 600))
sql = open(os.path.join(ROOT, "harness", "c1_query.sql")).read().replace("__THREADS__", str(a.threads)).replace("__SLICES__", "\n".join(slices))
env = dict(os.environ)
env["LD_LIBRARY_PATH"] = os.path.join(REFDIR, a.flavour) + ":" + env.get("LD_LIBRARY_PATH", "")
r = subprocess.run([os.path.join(REFDIR, "duckdb")], input=sql, capture_output=True, text=True, env=env, timeout=1800)
times = [float(x) for x in re.findall(r"Run Time \(s\): real ([0-9.]+)", r.stdout)]
per = 17  # slices of 600 groups
t1, t2 = sum(times[0:per]), sum(times[per + 1:2 * per + 1])
t3 = sum(times[2 * per + 1:3 * per + 1])
ok = len(times) == 3 * per + 1
res = {"flavour": a.flavour, "threads": a.threads,
       "query": "ols_fit_agg(y,[x1,x2,x3]) GROUP BY group_id over 10k groups x 100 rows, issued as 17 slices of 600 groups "
                "(the reference's Finalize overruns its LIST child vector beyond 2048/p groups per chunk)",
       "seconds_no_inference": [t1, t2] if ok else None, "seconds_with_inference": t3 if ok else None,
       "fits_per_sec_best": (10000 / min(t1, t2)) if ok else None,
       "stdout_tail": r.stdout[-600:], "stderr_tail": r.stderr[-300:]}
print(json.dumps(res))
if a.out:
    json.dump(res, open(a.out, "w"), indent=1)
