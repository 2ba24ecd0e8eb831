# A/B helper: engine options (and variant libraries under ab/) on a bench workload (WL=trustI|galaxy|magtar)
run() { lib=$1; shift; o="$*"; args=""; for kv in $o; do args="$args --option $kv"; done
  DARTRAY_ENGINE_LIB=$lib python bench.py --workload ${WL:-trustI} --no-also --no-cpu --no-e2e --steps 2 --warmup 1 $args 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('${WL:-trustI} $lib $o', round(d['ms_per_step'],1), round(d['roofline']['kernel_ms_per_step'],1), round(d['roofline']['frac'],4), d['gpu_launches'])"; }
for o in chunk=2 chunk=4 chunk=16 chunk=32; do run dart-ray_b200/libdartray_engine.so $o; WL=galaxy run dart-ray_b200/libdartray_engine.so $o; done
