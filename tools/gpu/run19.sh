set -x
cd $GRAFT_REPO_ROOT
(timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q) > gpurun_out/r02s_no_alloc_alone.log 2>&1
tail -5 gpurun_out/r02s_no_alloc_alone.log
(time timeout 2400 python -m pytest tests -m gpu -x -q) > gpurun_out/r02s_gpu_tests.log 2>&1
tail -8 gpurun_out/r02s_gpu_tests.log
