#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python tools/kernel_variants.py --only fresnel,teapot,dragons,csg,skybox --out gpurun_out/run7_variants.jsonl --reps 7 \
  r1:variants/r1.so: now:: 2>&1 | tee gpurun_out/run7_variants.log
