#!/bin/bash
# usage: tools/sweep.sh WORKLOAD "kb list" "team list"
w=$1
for kb in $2; do for tm in $3; do
  ANOFOX_B200_RING_WIN_KB=$kb ANOFOX_B200_RING_TEAM=$tm timeout 300 python bench.py --workload $w --no-cpu --no-e2e --steps 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('$w kb=$kb team=$tm acc_ms', round(r['avg_launch_ms'],3), 'frac', round(r['frac'],3))"
done; done
