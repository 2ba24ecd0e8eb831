set -x
cd $GRAFT_REPO_ROOT
for nc in 0 1; do
if [ $nc = 1 ]; then export BENCH_NO_CLOCKS=1; fi
BENCH_DIAG=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2956$nc bench.py --gpus 8 --steps 12 --warmup 3 --no-cpu --no-e2e --no-also > gpurun_out/r02z_diag_8gpu_noclk$nc.json 2> gpurun_out/r02z_diag_8gpu_noclk$nc.err
tail -c 150 gpurun_out/r02z_diag_8gpu_noclk$nc.json
done
