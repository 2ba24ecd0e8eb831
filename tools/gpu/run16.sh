set -x
cd $GRAFT_REPO_ROOT
wl=galaxy
CMD="python bench.py --workload $wl --steps 1 --warmup 1 --no-also --no-cpu --no-e2e"
$CMD > gpurun_out/r02p_plain_$wl.json 2> gpurun_out/r02p_plain_$wl.err &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02p_launches_$wl.csv $CMD > gpurun_out/r02p_ncu_$wl.log 2>&1
IDX=$(python tools/gpu/pick_launch.py gpurun_out/r02p_launches_$wl.csv k_deposit)
echo "longest k_deposit launch of $wl: $IDX"
$CMD > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_deposit -s $IDX -c 1 -o gpurun_out/r02p_kdeposit_$wl $CMD > gpurun_out/r02p_ncu_full_$wl.log 2>&1
tail -3 gpurun_out/r02p_ncu_full_$wl.log
