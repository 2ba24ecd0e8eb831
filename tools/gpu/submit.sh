#!/bin/bash
# usage: tools/gpu/submit.sh <script under tools/gpu> <timeout_s> [gpus]   -- retries while the pod answers busy (rc 3)
s=$1; t=${2:-1800}; g=${3:-1}
for i in $(seq 1 40); do
  if [ "$g" = "1" ]; then gpurun --timeout $t -- "bash tools/gpu/$s" > gpurun_out/$s.out 2>&1; else gpurun --gpus $g --timeout $t -- "bash tools/gpu/$s" > gpurun_out/$s.out 2>&1; fi
  rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" gpurun_out/$s.out; then break; fi
  sleep 90
done
echo "submit done rc=$rc" >> gpurun_out/$s.out
