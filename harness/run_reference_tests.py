#!/usr/bin/env python3
"""Run the reference's own sqllogictest files for the hot path through the drop-in harness.

The harness (harness/duckdb, built by oracle/build_ref.sh) is DuckDB + the reference's UNMODIFIED aggregate / window /
table-function sources for OLS, Ridge and WLS, linked against "libanofox_ffi.so". LD_LIBRARY_PATH picks the flavour:
  cpu   oracle/_ref/cpu   -> the CPU oracle behind the seven reference symbols
  b200  oracle/_ref/b200  -> libanofox_stats_b200.so (the drop-in proof; needs a B200)

prepare (build container, /root/reference present): every listed .test file is run with the cpu flavour; a record whose
ONLY problem is that it calls a function of another model family (Catalog Error "... does not exist": the trimmed
extension registers the hot path only) is dropped and the file is run again, until it passes or fails for a real
reason. The surviving files are written to oracle/_ref/tests/ (git-ignored build output that travels to the GPU box).
run (GPU box or here): the prepared files are run as they are with the chosen flavour.

usage: run_reference_tests.py prepare|run [--flavour cpu|b200] [--out file.json]
"""
import argparse
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
REFDIR = os.path.join(ROOT, "oracle", "_ref")
FILES = [
    "test/sql/ols_validation.test",
    "test/sql/regression/test_fit_agg.test",
    "test/sql/regression/test_map_options.test",
    "test/sql/fit_predict_agg/test_ols_fit_predict_agg.test",
    "test/sql/fit_predict_agg/test_ridge_fit_predict_agg.test",
    "test/sql/fit_predict_agg/test_wls_fit_predict_agg.test",
    "test/sql/fit_predict/test_ols_fit_predict_basic.test",
    "test/sql/fit_predict/test_ridge_fit_predict_basic.test",
    "test/sql/fit_predict/test_ridge_fit_predict_edge.test",
    "test/sql/fit_predict/test_ridge_fit_predict_rolling.test",
    "test/sql/fit_predict/test_ridge_fit_predict_validation.test",
    "test/sql/fit_predict/test_wls_fit_predict_basic.test",
    "test/sql/fit_predict/test_wls_fit_predict_edge.test",
    "test/sql/fit_predict/test_wls_fit_predict_rolling.test",
    "test/sql/fit_predict/test_wls_fit_predict_validation.test",
    "test/sql/macros/test_fit_predict_by.test",
    "test/sql/anofox_stats.test",
    "test/sql/comprehensive_tests.test",
    "test/sql/error_handling.test",
    "test/sql/type_handling.test",
]


def run_unittest(test_root, rel, flavour):
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = os.path.join(REFDIR, flavour) + ":" + env.get("LD_LIBRARY_PATH", "")
    exe = os.path.join(REFDIR, "unittest")
    r = subprocess.run([exe, "--test-dir", test_root, rel], capture_output=True, text=True, env=env, timeout=600)
    out = r.stdout + r.stderr
    m = re.search(r"assertions:\s*(\d+)\s*\|\s*(\d+) passed(?:\s*\|\s*(\d+) failed)?", out)
    allp = re.search(r"All tests passed \((\d+) assertions?", out)
    if allp:
        return True, int(allp.group(1)), out
    passed = int(m.group(2)) if m else 0
    return False, passed, out


def split_records(text):
    """sqllogictest records are separated by blank lines; keep the separators so that the file can be re-joined."""
    recs, cur = [], []
    for ln in text.split("\n"):
        if ln.strip() == "":
            if cur:
                recs.append(cur)
                cur = []
        else:
            cur.append(ln)
    if cur:
        recs.append(cur)
    return recs


def failing_line(out, rel):
    m = re.search(re.escape(rel) + r":(\d+)\)!", out)
    return int(m.group(1)) if m else None


def missing_function(out):
    m = re.search(r"(?:Catalog Error|Binder Error)[^\n]*?(?:Function|Macro|Table Function|Table)[^\n]*? with name \"?([A-Za-z0-9_]+)\"? does not exist", out)
    if m:
        return m.group(1)
    m = re.search(r"with name ([A-Za-z0-9_]+) does not exist", out)
    return m.group(1) if m else None


HOT = re.compile(r"^(anofox_stats_)?(ols|ridge|wls)_(fit|fit_agg|fit_predict|fit_predict_agg|fit_predict_by|predict_agg)$")


def prepare(args):
    os.makedirs(os.path.join(REFDIR, "tests"), exist_ok=True)
    summary = {}
    for rel in FILES:
        src = os.path.join(REF, rel)
        if not os.path.exists(src):
            continue
        text = open(src).read()
        recs = split_records(text)
        dropped = []
        stage = os.path.join(REFDIR, "tests")
        dst = os.path.join(stage, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        result = None
        for _ in range(len(recs) + 1):
            open(dst, "w").write("\n\n".join("\n".join(r) for r in recs) + "\n")
            ok, passed, out = run_unittest(stage, rel, "cpu")
            if ok:
                result = dict(status="pass", assertions=passed)
                break
            fn = missing_function(out)
            ln = failing_line(out, rel)
            if fn is None or HOT.match(fn) or ln is None:
                tail = out[out.find("=====") if "=====" in out else 0:][:1500]
                result = dict(status="fail", assertions_passed=passed, line=ln, detail=tail)
                break
            # drop the record that holds line ln of the CURRENT file
            cur, hit = 1, None
            for i, r in enumerate(recs):
                if cur <= ln < cur + len(r) + 1:
                    hit = i
                    break
                cur += len(r) + 1
            if hit is None:
                result = dict(status="fail", assertions_passed=passed, line=ln, detail="could not locate the record")
                break
            dropped.append(dict(function=fn, first_line=recs[hit][0][:100]))
            del recs[hit]
        result["records_kept"] = len(recs)
        result["records_dropped_other_families"] = len(dropped)
        result["dropped_functions"] = sorted({d["function"] for d in dropped})
        summary[rel] = result
        print(rel, result["status"], "kept", len(recs), "dropped", len(dropped), result.get("assertions", result.get("assertions_passed")))
        if result["status"] != "pass":
            print(result.get("detail", "")[:800])
    json.dump(summary, open(args.out, "w"), indent=1, sort_keys=True)


def run(args):
    stage = os.path.join(REFDIR, "tests")
    summary = {}
    for rel in FILES:
        if not os.path.exists(os.path.join(stage, rel)):
            continue
        ok, passed, out = run_unittest(stage, rel, args.flavour)
        summary[rel] = dict(status="pass" if ok else "fail", assertions=passed)
        if not ok:
            summary[rel]["detail"] = out[out.find("====") if "====" in out else 0:][:1500]
        print(rel, summary[rel]["status"], passed)
    summary["_flavour"] = args.flavour
    json.dump(summary, open(args.out, "w"), indent=1, sort_keys=True)
    return 0 if all(v.get("status") == "pass" for k, v in summary.items() if not k.startswith("_")) else 1


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("mode", choices=["prepare", "run"])
    ap.add_argument("--flavour", default="cpu", choices=["cpu", "b200"])
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "reference_tests.json"))
    a = ap.parse_args()
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    sys.exit(prepare(a) if a.mode == "prepare" else run(a))
