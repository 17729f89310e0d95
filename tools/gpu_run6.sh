#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/run6_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run6_pytest.log
tail -4 gpurun_out/run6_pytest.log
timeout 900 python tools/kernel_variants.py --only fresnel,cover,teapot,dragons,csg,skybox --out gpurun_out/run6_variants.jsonl --reps 7 \
  r1:variants/r1.so: now:: nolean::RTC_LEAN=0 2>&1 | tee gpurun_out/run6_variants.log
