set -x
cd $GRAFT_REPO_ROOT
(time timeout 2400 python -m pytest tests -m gpu -x -q) > gpurun_out/r02w_gpu_tests.log 2>&1
tail -8 gpurun_out/r02w_gpu_tests.log
