set -x
cd $GRAFT_REPO_ROOT
(time timeout 2400 python -m pytest tests -m gpu -x -q) > gpurun_out/r02h_gpu_tests.log 2>&1
tail -15 gpurun_out/r02h_gpu_tests.log
DRE_VERBOSE=1 timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02h_bench_trustI.json 2> gpurun_out/r02h_bench_trustI.err
tail -c 600 gpurun_out/r02h_bench_trustI.err
