#!/bin/bash
cd "$(dirname "$0")/.."
for c in 0 4 8 16 32; do RTC_COPY_CTAS=$c timeout 120 python tools/e2e_probe.py dragons 1920 1080 2>&1 | tail -1; done | tee gpurun_out/run22_e2e.log
for c in 0 16; do for sc in "cover 1920 1080" "teapot 1920 1080" "skybox 3840 2160"; do RTC_COPY_CTAS=$c timeout 120 python tools/e2e_probe.py $sc 2>&1 | tail -1; done; done | tee -a gpurun_out/run22_e2e.log
