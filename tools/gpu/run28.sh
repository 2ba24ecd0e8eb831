set -x
cd $GRAFT_REPO_ROOT
for i in 1 2 3; do
pw=1.5; if [ $i = 2 ]; then pw=0; fi
BENCH_DIAG=1 timeout 300 python bench.py --steps 16 --warmup 3 --prewarm-s $pw --no-cpu --no-e2e --no-also > gpurun_out/r02zx_diag_$i.json 2> gpurun_out/r02zx_diag_$i.err
tail -c 100 gpurun_out/r02zx_diag_$i.json
done
