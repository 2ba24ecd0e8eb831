set -x
cd $GRAFT_REPO_ROOT
(time timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_zedge.py -m gpu -x -q) > gpurun_out/r02j_gpu_tests.log 2>&1
tail -6 gpurun_out/r02j_gpu_tests.log
DRE_VERBOSE=1 timeout 1200 python bench.py --steps 5 --warmup 3 > gpurun_out/r02j_bench_trustI.json 2> gpurun_out/r02j_bench_trustI.err
tail -c 600 gpurun_out/r02j_bench_trustI.err
