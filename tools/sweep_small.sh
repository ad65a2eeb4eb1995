#!/bin/bash
# usage: tools/sweep_small.sh WORKLOAD "rounds list" "kb list"   -- stage geometry of the small-group kernel
w=$1
for rd in $2; do for kb in $3; do
  ANOFOX_B200_SMALL_ROUNDS=$rd ANOFOX_B200_SMALL_KB=$kb timeout 300 python bench.py --workload $w --no-cpu --no-e2e --steps 20 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('$w rounds=$rd kb=$kb step_ms', round(d['ms_per_step'],4), 'acc_ms', round(r['avg_launch_ms'],4), 'frac', round(r['frac'],3), 'ok', d.get('groups_fitted_ok'))"
done; done
