set -x
cd $GRAFT_REPO_ROOT
for b in 0 1; do
timeout 60 python bench.py --steps 120 --warmup 3 --prewarm-s 0.5 --no-cpu --no-e2e --no-also --option blocking_sync=$b > gpurun_out/r02zs_blocking$b.json 2> gpurun_out/r02zs_blocking$b.err
tail -c 100 gpurun_out/r02zs_blocking$b.json
done
