#!/bin/bash
cd "$(dirname "$0")/.."
for c in 0 4 8 16; do RTC_COPY_CTAS=$c timeout 120 python tools/e2e_probe.py dragons 1920 1080 2>&1 | tail -1; done | tee gpurun_out/run29_e2e.log
for c in 0 8; do for sc in "cover 1920 1080" "teapot 1920 1080" "skybox 3840 2160" "csg 3840 2160" "fresnel 600 600"; do RTC_KERNEL=mega RTC_COPY_CTAS=$c timeout 120 python tools/e2e_probe.py $sc 2>&1 | tail -1; done; done | tee -a gpurun_out/run29_e2e.log
timeout 600 python -m pytest tests -m gpu -x -q -k "smoke or partition or rgba8 or ragged or baseline_config" > gpurun_out/run29_pytest.log 2>&1; tail -3 gpurun_out/run29_pytest.log
