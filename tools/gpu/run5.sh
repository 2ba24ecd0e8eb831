set -x
cd $GRAFT_REPO_ROOT
(time timeout 1700 python -m pytest tests -m gpu -x -q) > gpurun_out/r02e_gpu_tests.log 2>&1
tail -15 gpurun_out/r02e_gpu_tests.log
DRE_VERBOSE=2 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/r02e_bench_trustI.json 2> gpurun_out/r02e_bench_trustI.err
tail -c 600 gpurun_out/r02e_bench_trustI.err
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-also --no-e2e --option coop=0 > gpurun_out/r02e_bench_trustI_nocoop.json 2> gpurun_out/r02e_bench_trustI_nocoop.err
DARTRAY_ENGINE_LIB=ab/libdre_unroll2.so timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/r02e_bench_trustI_unroll2.json 2> gpurun_out/r02e_bench_trustI_unroll2.err
