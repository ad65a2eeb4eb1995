#!/usr/bin/env python3
"""Summarise an .ncu-rep here (no GPU): key raw metrics + instruction / stall-sample shares per source line.
usage: tools/ncu_summary.py report.ncu-rep [top_n] [kernel-name-regex]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
ksel = ["--kernel-name", "regex:" + sys.argv[3]] if len(sys.argv) > 3 else []
raw = subprocess.run(["ncu", "-i", rep] + ksel + ["--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
print("kernel:", vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?")
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "sm__cycles_elapsed.max", "sass__inst_executed_local_loads", "sass__inst_executed_local_stores",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"]
for h, u, v in zip(hdr, units, vals):
    if h in want or ("pcsamp_warps_issue_stalled" in h and "not_issued" not in h):
        print(f"{h:90s} {u:12s} {v}")
src = subprocess.run(["ncu", "-i", rep] + ksel + ["--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
agg, fileinfo = {}, None
for r in csv.reader(io.StringIO(src)):
    if len(r) >= 2 and r[0] == "File Path":
        fileinfo = r[1].split("/")[-1]
        continue
    if len(r) < 8 or not r[0] or r[0] == "Line No" or r[2] != "-":
        continue
    try:
        key = (fileinfo, int(r[0]), r[1].strip()[:80])
        a = agg.get(key, (0, 0))
        agg[key] = (a[0] + int(r[7]), a[1] + int(r[4]))
    except ValueError:
        pass
tot = sum(v[0] for v in agg.values()) or 1
ts = sum(v[1] for v in agg.values()) or 1
print(f"\nwarp instructions {tot}, stall samples {ts}; top {top} source lines by instructions:")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{v[0] / 1e6:9.2f}M {100 * v[0] / tot:5.1f}%  samples {100 * v[1] / ts:5.1f}%  {k[0]}:{k[1]}  {k[2]}")
print(f"\ntop {top} source lines by stall samples:")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{v[0] / 1e6:9.2f}M {100 * v[0] / tot:5.1f}%  samples {100 * v[1] / ts:5.1f}%  {k[0]}:{k[1]}  {k[2]}")

# SASS view: which instructions the warps wait at (top by stall samples, with the opcode)
sass = subprocess.run(["ncu", "-i", rep] + ksel + ["--page", "source", "--print-source", "sass", "--csv"], capture_output=True, text=True).stdout
rows = [r for r in csv.reader(io.StringIO(sass))]
hdr_i = next((i for i, r in enumerate(rows) if "Source" in r and any("Samples" in c for c in r)), None)
if hdr_i is not None:
    h = rows[hdr_i]
    ci = h.index("Source")
    si = next(i for i, c in enumerate(h) if "Samples" in c)
    body = []
    for r in rows[hdr_i + 1:]:
        try:
            body.append((int(r[si]), r[ci].strip()[:110]))
        except (ValueError, IndexError):
            pass
    tot_s = sum(b[0] for b in body) or 1
    print(f"\ncolumns: {h[:12]}\ntop 40 SASS instructions by '{h[si]}' (total {tot_s}):")
    for n, (smp, txt) in sorted(enumerate(body), key=lambda kv: -kv[1][0])[:40]:
        print(f"  #{n:6d} {100 * smp / tot_s:5.1f}%  {txt}")
    # by opcode
    byop = {}
    for smp, txt in body:
        t = txt.split()
        op = next((w for w in t if not w.startswith("@") and w[0].isalpha()), "?").split(".")[0]
        byop[op] = byop.get(op, 0) + smp
    print("stall samples by opcode:", ", ".join(f"{k} {100 * v / tot_s:.1f}%" for k, v in sorted(byop.items(), key=lambda kv: -kv[1])[:16]))
