#!/bin/bash
# round-2 GPU call 1: correctness of the new build, kernel variants, scaling probe, ncu capture
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > gpurun_out/run1_gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/run1_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run1_pytest.log
tail -3 gpurun_out/run1_pytest.log
rm -f gpurun_out/variants.jsonl
timeout 900 python tools/kernel_variants.py --only dragons,teapot,cover \
  r1:variants/r1.so: new:: noinst::RTC_INSTANCING=0 nolean::RTC_LEAN=0 noearly:variants/noearly.so: \
  leaf2::RTC_BVH_LEAF=2 leaf3::RTC_BVH_LEAF=3 leaf6::RTC_BVH_LEAF=6 2>&1 | tee gpurun_out/run1_variants.log
timeout 300 python tools/kernel_variants.py --tool scaling_probe --only dragons --out gpurun_out/run1_probe.jsonl new:: 2>&1 | tee gpurun_out/run1_probe.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 5 -c 1 -f -o gpurun_out/r2_a_dragons \
  python tools/bench_all.py --only dragons --reps 6 > gpurun_out/run1_ncu.log 2>&1
echo done
