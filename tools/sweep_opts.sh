# A/B: threads per CTA of the ray-march kernel (variant libraries under ab/) x stage_split
run() { lib=$1; shift; o="$*"; args=""; for kv in $o; do args="$args --option $kv"; done
  DARTRAY_ENGINE_LIB=$lib python bench.py --no-cpu --no-e2e --steps 2 --warmup 1 $args 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$lib $o', round(d['ms_per_step'],1), round(d['roofline']['kernel_ms_per_step'],1))"; }
run dart-ray_b200/libdartray_engine.so stage_split=2 warps=16
run ab/eng_640.so stage_split=4 warps=20
run ab/eng_640.so stage_split=8 warps=20
run ab/eng_640.so stage_split=4 warps=18
run ab/eng_768.so stage_split=8 warps=24
run ab/eng_768.so stage_split=8 warps=22
run ab/eng_768.so stage_split=4 warps=20
