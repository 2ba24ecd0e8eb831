set -x
cd $GRAFT_REPO_ROOT
(time timeout 900 python -m pytest tests/test_gpu_abi_host.py tests/test_gpu_dropin.py tests/test_gpu_parity.py -m gpu -x -q) > gpurun_out/r02d_gpu_tests.log 2>&1
tail -5 gpurun_out/r02d_gpu_tests.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02d_bench_trustI.json 2> gpurun_out/r02d_bench_trustI.err
tail -c 800 gpurun_out/r02d_bench_trustI.err
for wl in trustI magtar; do
CMD="python bench.py --workload $wl --steps 1 --warmup 1 --no-also --no-cpu --no-e2e"
$CMD > gpurun_out/r02d_plain_$wl.json 2> gpurun_out/r02d_plain_$wl.err &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02d_launches_$wl.csv $CMD > gpurun_out/r02d_ncu_$wl.log 2>&1
IDX=$(python tools/gpu/pick_launch.py gpurun_out/r02d_launches_$wl.csv k_deposit)
echo "longest k_deposit launch of $wl: $IDX"
$CMD > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_deposit -s $IDX -c 1 -o gpurun_out/r02d_kdeposit_$wl $CMD > gpurun_out/r02d_ncu_full_$wl.log 2>&1
done
ls -la gpurun_out/
