set -x
cd $GRAFT_REPO_ROOT
DRE_VERBOSE=1 timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu --also galaxy,magtar > gpurun_out/r02u_bench_trustI.json 2> gpurun_out/r02u_bench_trustI.err
grep -c refused gpurun_out/r02u_bench_trustI.err
tail -c 200 gpurun_out/r02u_bench_trustI.json
