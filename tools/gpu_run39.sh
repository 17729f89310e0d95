#!/bin/bash
cd "$(dirname "$0")/.."
for v in "RTC_DMA_STREAMS=1" "RTC_DMA_STREAMS=2" "RTC_DMA_STREAMS=4" "RTC_DMA_STREAMS=2 RTC_DMA_MIN_KB=512" "RTC_DMA_STREAMS=4 RTC_DMA_MIN_KB=256" "RTC_DMA_STREAMS=2 RTC_DMA_MIN_KB=2048" "RTC_DMA_STREAMS=2 RTC_COPY_BUCKETS=0" "RTC_DMA_STREAMS=2 RTC_COPY_BUCKETS=2"; do
  echo "== $v"; env $v timeout 300 python tools/e2e_probe.py dragons 1920 1080 2>&1 | tail -2
done | tee gpurun_out/run39_e2e.log
