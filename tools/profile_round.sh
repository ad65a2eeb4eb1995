#!/bin/bash
# usage: tools/profile_round.sh <tag>   -- the ncu evidence of a round (run under gpurun, one GPU)
# launch list of the default bench command, then one --set full capture of the top kernels of C2, C3r, C5.
tag=${1:-r01}
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e"
$B > gpurun_out/plain_c2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_c2.csv $B > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"fit_small|solve_small" -s 6 -c 2 -o gpurun_out/${tag}_prof_c2 $B > gpurun_out/ncu2.log 2>&1
$B --workload C3r > gpurun_out/plain_c3r.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"acc_ring|solve_fixed" -s 6 -c 2 -o gpurun_out/${tag}_prof_c3r $B --workload C3r > gpurun_out/ncu3.log 2>&1
$B --workload C5 > gpurun_out/plain_c5.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"acc_ring|predict_flat" -s 4 -c 2 -o gpurun_out/${tag}_prof_c5 $B --workload C5 > gpurun_out/ncu4.log 2>&1
tail -2 gpurun_out/ncu1.log gpurun_out/ncu2.log gpurun_out/ncu3.log gpurun_out/ncu4.log | cut -c1-300
