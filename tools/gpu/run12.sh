set -x
cd $GRAFT_REPO_ROOT
(time timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_zedge.py tests/test_mode_2d.py -m gpu -x -q) > gpurun_out/r02l_gpu_tests.log 2>&1
tail -6 gpurun_out/r02l_gpu_tests.log
DRE_VERBOSE=2 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e --also galaxy,magtar > gpurun_out/r02l_bench_trustI.json 2> gpurun_out/r02l_bench_trustI.err
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e --also galaxy,magtar --option coop=0 > gpurun_out/r02l_bench_trustI_nocoop.json 2> gpurun_out/r02l_bench_trustI_nocoop.err
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e --no-also --option coop_work=1000000 > gpurun_out/r02l_bench_trustI_coop1m.json 2> gpurun_out/r02l_bench_trustI_coop1m.err
