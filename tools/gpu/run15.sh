set -x
cd $GRAFT_REPO_ROOT
run() { lib=$1; tag=$2; shift 2; args=""; for kv in "$@"; do args="$args --option $kv"; done
  DARTRAY_ENGINE_LIB=$lib timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e --also galaxy,magtar $args > gpurun_out/r02o_bench_$tag.json 2> gpurun_out/r02o_bench_$tag.err; }
run ab/libdre_t640.so t640_w20_s8 warps=20 stage_split=8
run ab/libdre_t640.so t640_w20_s6 warps=20 stage_split=6
run ab/libdre_t768.so t768_w24_s8 warps=24 stage_split=8
run dart-ray_b200/libdartray_engine.so t512_w16_s4 stage_split=4
