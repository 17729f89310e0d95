#!/bin/bash
# after the release-atomic fix (no L1 invalidation per tile): dma / warps modes, order buckets, run sizes
cd "$(dirname "$0")/.."
for v in "RTC_HOST_COPY=dma" "RTC_HOST_COPY=dma RTC_COPY_BUCKETS=0" "RTC_HOST_COPY=dma RTC_COPY_BUCKETS=2" "RTC_HOST_COPY=dma RTC_COPY_BUCKETS=3" "RTC_HOST_COPY=dma RTC_DMA_SPREAD=1" "RTC_HOST_COPY=dma RTC_DMA_MIN_KB=512" "RTC_HOST_COPY=dma RTC_DMA_MIN_KB=2048" "RTC_HOST_COPY=warps"; do
  echo "== $v"; env $v timeout 300 python tools/e2e_probe.py dragons 1920 1080 2>&1 | tail -2
done | tee gpurun_out/run38_e2e.log
for sc in "teapot 1920 1080" "skybox 3840 2160" "cover 1920 1080" "csg 3840 2160" "fresnel 600 600"; do
  for m in dma warps; do RTC_HOST_COPY=$m timeout 300 python tools/e2e_probe.py $sc 2>&1 | tail -2; done
done | tee -a gpurun_out/run38_e2e.log
