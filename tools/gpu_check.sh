#!/bin/bash
# usage: tools/gpu_check.sh [workloads...]  -- GPU tests, then a short bench line per workload
timeout 1200 python -m pytest tests -m gpu -x -q --timeout 300 > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
for w in "$@"; do
  timeout 300 python bench.py --workload $w --no-cpu --no-e2e --steps 5 > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_$w.json"))
    r = d["roofline"]
    print("$w", "step_ms", round(d["ms_per_step"], 3), "acc_ms", round(r["avg_launch_ms"], 3), "frac", round(r["frac"], 3),
          "solve_ms", round(r.get("solve_stage_ms", 0) or 0, 3), "ok", d.get("groups_fitted_ok"))
except Exception as e:
    print("$w failed", e)
PY
  tail -3 gpurun_out/bench_$w.err
done
