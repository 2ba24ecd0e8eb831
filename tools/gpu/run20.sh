set -x
cd $GRAFT_REPO_ROOT
nvidia-smi -L | wc -l
DRE_VERBOSE=2 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r02x_bench_trustI_8gpu.json 2> gpurun_out/r02x_bench_trustI_8gpu.err
tail -c 300 gpurun_out/r02x_bench_trustI_8gpu.json
