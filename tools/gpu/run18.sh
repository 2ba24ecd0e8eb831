set -x
cd $GRAFT_REPO_ROOT
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu --also galaxy,magtar > gpurun_out/r02r_bench_trustI.json 2> gpurun_out/r02r_bench_trustI.err
tail -c 300 gpurun_out/r02r_bench_trustI.err
(time timeout 2400 python -m pytest tests -m gpu -x -q) > gpurun_out/r02r_gpu_tests.log 2>&1
tail -8 gpurun_out/r02r_gpu_tests.log
