set -x
cd $GRAFT_REPO_ROOT
for v in dart-ray_b200/libdartray_engine.so ab/libdre_g6.so ab/libdre_g8.so; do
n=$(basename $v .so)
DARTRAY_ENGINE_LIB=$v timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e --also magtar > gpurun_out/r02n_bench_$n.json 2> gpurun_out/r02n_bench_$n.err
done
timeout 1500 python bench.py --workload galaxy_full --steps 1 --warmup 1 --no-also --no-cpu --no-e2e > gpurun_out/r02n_bench_galaxy_full.json 2> gpurun_out/r02n_bench_galaxy_full.err
tail -c 300 gpurun_out/r02n_bench_galaxy_full.err
