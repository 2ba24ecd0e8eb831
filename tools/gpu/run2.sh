set -x
cd $GRAFT_REPO_ROOT
(time timeout 1700 python -m pytest tests -m gpu -x -q) > gpurun_out/r02b_gpu_tests.log 2>&1
tail -15 gpurun_out/r02b_gpu_tests.log
DRE_VERBOSE=1 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/r02b_bench_trustI.json 2> gpurun_out/r02b_bench_trustI.err
tail -c 1500 gpurun_out/r02b_bench_trustI.err
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-also --no-e2e --option graphs=0 > gpurun_out/r02b_bench_trustI_nographs.json 2> gpurun_out/r02b_bench_trustI_nographs.err
