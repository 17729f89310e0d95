#!/bin/bash
# host canvases through the copy engine (RTC_HOST_COPY=dma) against copy warps and direct stores; GPU tests first
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/run37_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run37_pytest.log
tail -4 gpurun_out/run37_pytest.log
for mode in dma warps direct; do
  for sc in "dragons 1920 1080" "teapot 1920 1080" "skybox 3840 2160" "cover 1920 1080"; do
    RTC_HOST_COPY=$mode timeout 300 python tools/e2e_probe.py $sc 2>&1 | tail -2
  done
done | tee gpurun_out/run37_e2e.log
for v in "RTC_DMA_SPREAD=1" "RTC_COPY_BUCKETS=0" "RTC_COPY_BUCKETS=3" "RTC_DMA_MIN_KB=256" "RTC_DMA_MIN_KB=4096"; do
  echo "== $v"; env $v timeout 300 python tools/e2e_probe.py dragons 1920 1080 2>&1 | tail -2
done | tee -a gpurun_out/run37_e2e.log
