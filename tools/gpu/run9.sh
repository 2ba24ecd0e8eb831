set -x
cd $GRAFT_REPO_ROOT
DRE_VERBOSE=1 timeout 1200 python bench.py --steps 5 --warmup 3 > gpurun_out/r02i_bench_trustI.json 2> gpurun_out/r02i_bench_trustI.err
tail -c 600 gpurun_out/r02i_bench_trustI.err
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02i_smoke.log 2>&1; tail -2 gpurun_out/r02i_smoke.log
