set -x
cd $GRAFT_REPO_ROOT
nvidia-smi -L
(time timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_zedge.py -m gpu -x -q -k "two_gpus or two_gpu or multigpu or 2gpu or ranks") > gpurun_out/r02zw_gpu_tests_2gpu.log 2>&1
tail -8 gpurun_out/r02zw_gpu_tests_2gpu.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu --also galaxy,magtar > gpurun_out/r02zw_bench_trustI_2gpu.json 2> gpurun_out/r02zw_bench_trustI_2gpu.err
tail -c 300 gpurun_out/r02zw_bench_trustI_2gpu.json
