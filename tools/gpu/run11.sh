set -x
cd $GRAFT_REPO_ROOT
(time timeout 900 python -m pytest tests/test_mode_2d.py tests/test_abi.py tests/test_gpu_abi_host.py tests/test_gpu_dropin.py -m gpu -x -q) > gpurun_out/r02k_gpu_tests.log 2>&1
tail -30 gpurun_out/r02k_gpu_tests.log
