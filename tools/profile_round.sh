#!/bin/bash
# usage: tools/profile_round.sh <tag>   -- the ncu evidence of a round (run under gpurun, one GPU)
# launch list of the default bench command, then one --set full capture of the top kernels of C2, C3r, C3w, C5, W1.
# The reports are summarised on the box (tools/ncu_summary.py, tools/ncu_traffic.py) and deleted: gpurun_out/ is
# limited to 64 MiB.
tag=${1:-r02}
O=gpurun_out
export TRAFFIC_JSON=$O/traffic.json
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-secondary"
$B > $O/plain_c2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${tag}_launches_c2.csv $B > $O/ncu1.log 2>&1
cap() { # name workload kernel-regex skip count
	$B --workload $2 > $O/plain_$1.log 2>&1 || return
	ncu --set full --clock-control none --import-source on -k regex:"$3" -s $4 -c $5 -o $O/${tag}_prof_$1 $B --workload $2 > $O/ncu_$1.log 2>&1
}
summ() { # name kernel-regex out-name [traffic-workload]
	python tools/ncu_summary.py $O/${tag}_prof_$1.ncu-rep 30 "$2" > $O/${tag}_$3_summary.txt 2>&1
	if [ -n "$4" ]; then python tools/ncu_traffic.py $4 $O/${tag}_prof_$1.ncu-rep "$2" ${tag} >> $O/traffic.log 2>&1; fi
}
cap c2 C2 "fit_small_kernel|solve_small" 6 2
summ c2 fit_small_kernel c2_fit_small C2; summ c2 solve_small c2_solve_small
rm -f $O/${tag}_prof_c2.ncu-rep
cap c3r C3r "acc_ring|solve_fixed" 6 2
summ c3r acc_ring c3r_acc_ring C3r; summ c3r solve_fixed c3r_solve_fixed
rm -f $O/${tag}_prof_c3r.ncu-rep
cap c3w C3w "acc_ring" 3 1
summ c3w acc_ring c3w_acc_ring C3w
rm -f $O/${tag}_prof_c3w.ncu-rep
cap c5 C5 "acc_ring|predict_flat" 4 2
summ c5 acc_ring c5_acc_ring C5; summ c5 predict_flat c5_predict_flat
rm -f $O/${tag}_prof_c5.ncu-rep
cap w1 W1 "window_fit_predict" 3 1
summ w1 window_fit_predict w1_window W1
rm -f $O/${tag}_prof_w1.ncu-rep
tail -n 2 $O/ncu1.log $O/ncu_c2.log $O/ncu_c3r.log $O/ncu_c5.log $O/ncu_w1.log | cut -c1-200
cat $O/traffic.log
