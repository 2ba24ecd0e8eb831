set -x
cd $GRAFT_REPO_ROOT
DRE_VERBOSE=2 timeout 600 python bench.py --workload trustI --steps 1 --warmup 1 --no-also --no-cpu --no-e2e > gpurun_out/r02zt_verbose_trustI.json 2> gpurun_out/r02zt_step_timeline_trustI.txt
tail -c 300 gpurun_out/r02zt_verbose_trustI.json
DRE_VERBOSE=2 timeout 600 python bench.py --workload magtar --steps 1 --warmup 1 --no-also --no-cpu --no-e2e > gpurun_out/r02zt_verbose_magtar.json 2> gpurun_out/r02zt_step_timeline_magtar.txt
grep -c "dre step" gpurun_out/r02zt_step_timeline_*.txt
