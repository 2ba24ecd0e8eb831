set -x
cd $GRAFT_REPO_ROOT
(time timeout 2400 python -m pytest tests -m gpu -x -q) > gpurun_out/r02m_gpu_tests.log 2>&1
tail -8 gpurun_out/r02m_gpu_tests.log
timeout 1200 python bench.py --steps 5 --warmup 3 > gpurun_out/r02m_bench_trustI.json 2> gpurun_out/r02m_bench_trustI.err
tail -c 500 gpurun_out/r02m_bench_trustI.err
