set -x
cd $GRAFT_REPO_ROOT
timeout 1200 python bench.py > gpurun_out/r02zz_bench_default.json 2> gpurun_out/r02zz_bench_default.err
tail -c 300 gpurun_out/r02zz_bench_default.err
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/r02zz_bench_reference.json 2> gpurun_out/r02zz_bench_reference.err
wl=trustI
CMD="python bench.py --workload $wl --steps 1 --warmup 1 --no-also --no-cpu --no-e2e"
$CMD > gpurun_out/r02zz_plain_$wl.json 2> gpurun_out/r02zz_plain_$wl.err &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02zz_launches_$wl.csv $CMD > gpurun_out/r02zz_ncu_$wl.log 2>&1
for k in k_prep_geom k_geom_write k_step_build; do
IDX=$(python tools/gpu/pick_launch.py gpurun_out/r02zz_launches_$wl.csv $k)
echo "longest $k launch of $wl: $IDX"
$CMD > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$k -s $IDX -c 1 -o gpurun_out/r02zz_${k}_$wl $CMD > gpurun_out/r02zz_ncu_full_${k}.log 2>&1
done
(time timeout 2400 python -m pytest tests -m gpu -x -q) > gpurun_out/r02zz_gpu_tests.log 2>&1
tail -8 gpurun_out/r02zz_gpu_tests.log
