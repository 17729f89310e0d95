#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python tools/kernel_variants.py --only dragons,teapot --out gpurun_out/run4_variants.jsonl --reps 5 \
  nosteal:variants/nosteal.so: s1i1:variants/s1i1.so: s1i16:variants/s1i16.so: s1i24:variants/s1i24.so: s3i16:variants/s3i16.so: s3i24:variants/s3i24.so: s3i28d2:variants/s3i28d2.so: 2>&1 | tee gpurun_out/run4_variants.log
for v in variants/nosteal.so variants/s1i16.so variants/s3i24.so variants/s3i28d2.so; do
  echo "lib=$v" | tee -a gpurun_out/run4_probe.log
  RTC_B200_LIB=${v:+$PWD/$v} timeout 300 python tools/scaling_probe.py --scene dragons --parts 8 2>&1 | tee -a gpurun_out/run4_probe.log
done
