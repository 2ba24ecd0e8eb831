set -x
cd $GRAFT_REPO_ROOT
(time timeout 1700 python -m pytest tests -m gpu -x -q) > gpurun_out/r02c_gpu_tests.log 2>&1
tail -15 gpurun_out/r02c_gpu_tests.log
for o in split=1 split=0; do
DRE_VERBOSE=2 timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu --no-e2e --option $o > gpurun_out/r02c_bench_trustI_$o.json 2> gpurun_out/r02c_bench_trustI_$o.err
done
timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu --no-also > gpurun_out/r02c_bench_trustI_e2e.json 2> gpurun_out/r02c_bench_trustI_e2e.err
