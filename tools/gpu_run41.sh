#!/bin/bash
cd "$(dirname "$0")/.."
for v in "RTC_HOST_COPY=dma" "RTC_HOST_COPY=dma RTC_DMA_STREAMS=1" "RTC_HOST_COPY=dma RTC_DMA_STREAMS=3" "RTC_HOST_COPY=dma RTC_DMA_MIN_KB=1024" "RTC_HOST_COPY=dma RTC_DMA_MIN_KB=64" "RTC_HOST_COPY=dma RTC_ROW_LEAD=70" "RTC_HOST_COPY=dma RTC_ROW_LEAD=140"; do
  echo "== $v"; env $v timeout 300 python tools/e2e_probe.py dragons 1920 1080 2>&1 | tail -2
done | tee gpurun_out/run41_e2e.log
for sc in "teapot 1920 1080" "skybox 3840 2160" "cover 1920 1080" "csg 3840 2160"; do
  RTC_HOST_COPY=dma timeout 300 python tools/e2e_probe.py $sc 2>&1 | tail -2
done | tee -a gpurun_out/run41_e2e.log
