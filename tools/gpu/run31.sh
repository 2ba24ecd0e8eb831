set -x
cd $GRAFT_REPO_ROOT
CMD="python tools/gpu/galaxy_full_sample.py"
(time $CMD) > gpurun_out/r02zu_plain_galaxy_full_sample.log 2>&1 || exit 1
tail -4 gpurun_out/r02zu_plain_galaxy_full_sample.log
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02zu_launches_galaxy_full_sample.csv $CMD > gpurun_out/r02zu_ncu.log 2>&1
IDX=$(python tools/gpu/pick_launch.py gpurun_out/r02zu_launches_galaxy_full_sample.csv k_deposit)
echo "longest k_deposit launch: $IDX"
timeout 200 ncu --set full --clock-control none --import-source on -k regex:k_deposit -s $IDX -c 1 --kill on -o gpurun_out/r02zu_kdeposit_galaxy_full $CMD > gpurun_out/r02zu_ncu_full.log 2>&1
tail -3 gpurun_out/r02zu_ncu_full.log
