#!/bin/bash
cd "$(dirname "$0")/.."
for b in 1 4; do for c in 8 12; do RTC_COPY_CTAS=$c RTC_COPY_BUCKETS=$b timeout 120 python tools/e2e_probe.py dragons 1920 1080 2>&1 | tail -1 | sed "s/^/buckets=$b /"; done; done | tee gpurun_out/run27_e2e.log
for sc in "cover 1920 1080" "teapot 1920 1080" "skybox 3840 2160" "csg 3840 2160" "fresnel 600 600"; do RTC_COPY_CTAS=8 timeout 120 python tools/e2e_probe.py $sc 2>&1 | tail -1; done | tee -a gpurun_out/run27_e2e.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/run27_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run27_pytest.log
tail -4 gpurun_out/run27_pytest.log
