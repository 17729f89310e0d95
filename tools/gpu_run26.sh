#!/bin/bash
cd "$(dirname "$0")/.."
for sp in 0 1; do for b in 0 1 3; do RTC_COPY_CTAS=8 RTC_COPY_SPREAD=$sp RTC_COPY_BUCKETS=$b timeout 120 python tools/e2e_probe.py dragons 1920 1080 2>&1 | tail -1 | sed "s/^/spread=$sp buckets=$b /"; done; done | tee gpurun_out/run26_e2e.log
for g in 1 2 8; do RTC_ROW_GROUP=$g RTC_COPY_CTAS=8 timeout 120 python tools/e2e_probe.py dragons 1920 1080 2>&1 | tail -1 | sed "s/^/spread=1 buckets=1 rowgroup=$g /"; done | tee -a gpurun_out/run26_e2e.log
for sc in "cover 1920 1080" "teapot 1920 1080" "skybox 3840 2160" "csg 3840 2160" "fresnel 600 600"; do RTC_COPY_CTAS=8 timeout 120 python tools/e2e_probe.py $sc 2>&1 | tail -1; done | tee -a gpurun_out/run26_e2e.log
