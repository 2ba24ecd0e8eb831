set -x
cd $GRAFT_REPO_ROOT
nvidia-smi -L
nproc
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r02a_gpu_tests.log 2>&1
tail -5 gpurun_out/r02a_gpu_tests.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02a_bench_trustI.json 2> gpurun_out/r02a_bench_trustI.err
tail -c 600 gpurun_out/r02a_bench_trustI.err
for wl in magtar trustI galaxy; do
DRE_VERBOSE=0 DARTRAY_ENGINE_LIB=ab/libdre_timing.so timeout 300 python bench.py --workload $wl --no-also --no-cpu --no-e2e --steps 1 --warmup 1 > gpurun_out/r02a_timing_$wl.json 2> gpurun_out/r02a_timing_$wl.err
done
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02a_bench_reference.json 2> gpurun_out/r02a_bench_reference.err
