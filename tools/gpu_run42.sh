#!/bin/bash
cd "$(dirname "$0")/.."
RTC_DMA_TRACE=1 RTC_HOST_COPY=dma timeout 300 python tools/e2e_probe.py dragons 1920 1080 2>&1 | grep -v "^$" | tail -6 | cut -c1-1800 | tee gpurun_out/run42.log
