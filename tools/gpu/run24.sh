set -x
cd $GRAFT_REPO_ROOT
nproc
for rot in 1 0; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2955$rot bench.py --gpus 8 --steps 8 --warmup 3 --no-cpu --no-e2e --also galaxy,magtar --option sector_rotation=$rot > gpurun_out/r02y_bench_trustI_8gpu_rot$rot.json 2> gpurun_out/r02y_bench_trustI_8gpu_rot$rot.err
tail -c 200 gpurun_out/r02y_bench_trustI_8gpu_rot$rot.json
done
