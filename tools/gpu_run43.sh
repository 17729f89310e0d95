#!/bin/bash
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/run43_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run43_pytest.log
tail -4 gpurun_out/run43_pytest.log
for v in "RTC_HOST_COPY=dma" "RTC_HOST_COPY=dma RTC_BAND_KB=400" "RTC_HOST_COPY=dma RTC_BAND_KB=3072" "RTC_HOST_COPY=dma RTC_ROW_LEAD=150" "RTC_HOST_COPY=dma RTC_ROW_LEAD=70" "RTC_HOST_COPY=dma RTC_DMA_MIN_KB=256" "RTC_HOST_COPY=warps" "RTC_HOST_COPY=warps RTC_BAND_KB=200"; do
  echo "== $v"; env $v timeout 300 python tools/e2e_probe.py dragons 1920 1080 2>&1 | tail -2
done | tee gpurun_out/run43_e2e.log
for sc in "teapot 1920 1080" "skybox 3840 2160" "cover 1920 1080"; do
  RTC_HOST_COPY=dma timeout 300 python tools/e2e_probe.py $sc 2>&1 | tail -2
done | tee -a gpurun_out/run43_e2e.log
