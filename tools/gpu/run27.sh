set -x
cd $GRAFT_REPO_ROOT
for nc in 0 1; do
if [ $nc = 1 ]; then export BENCH_NO_CLOCKS=1; fi
BENCH_DIAG=1 timeout 300 python bench.py --steps 16 --warmup 3 --no-cpu --no-e2e --no-also > gpurun_out/r02zy_diag_1gpu_noclk$nc.json 2> gpurun_out/r02zy_diag_1gpu_noclk$nc.err
tail -c 100 gpurun_out/r02zy_diag_1gpu_noclk$nc.json
done
