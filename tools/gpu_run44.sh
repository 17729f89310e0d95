#!/bin/bash
cd "$(dirname "$0")/.."
timeout 600 python tools/bench_all.py --reps 7 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except ValueError: continue
    print(d['scene'], d.get('kernel'), round(d['kernel_ms'], 4), round(d['Mrays_s']))
" | tee gpurun_out/run44.log
for v in "RTC_HOST_COPY=dma" "RTC_HOST_COPY=dma RTC_ROW_LEAD=70" "RTC_HOST_COPY=warps"; do
  echo "== $v"; env $v timeout 300 python tools/e2e_probe.py dragons 1920 1080 2>&1 | tail -2
done | tee -a gpurun_out/run44.log
for sc in "teapot 1920 1080" "skybox 3840 2160" "cover 1920 1080" "csg 3840 2160" "fresnel 600 600"; do
  RTC_HOST_COPY=dma timeout 300 python tools/e2e_probe.py $sc 2>&1 | tail -2
done | tee -a gpurun_out/run44.log
