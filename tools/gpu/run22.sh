set -x
cd $GRAFT_REPO_ROOT
nproc; cat /proc/cpuinfo | grep "model name" | head -1
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu --also galaxy,magtar > gpurun_out/r02v_bench_trustI.json 2> gpurun_out/r02v_bench_trustI.err
tail -c 200 gpurun_out/r02v_bench_trustI.json
DRE_VERBOSE=2 timeout 600 python bench.py --workload magtar --steps 1 --warmup 2 --no-also --no-cpu --no-e2e > gpurun_out/r02v_verbose_magtar.json 2> gpurun_out/r02v_step_timeline_magtar.txt
