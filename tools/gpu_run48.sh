#!/bin/bash
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests -m gpu -x -q -k "pinned or copy or zero_copy or static" 2>&1 | tail -2
timeout 300 python tools/bench_all.py --only dragons,teapot --reps 7 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except ValueError: continue
    print(d['scene'], round(d['kernel_ms'], 4))"
for v in "RTC_HOST_COPY=dma" "RTC_HOST_COPY=dma RTC_ROW_LEAD=70" "RTC_HOST_COPY=dma RTC_COPY_BUCKETS=3" "RTC_HOST_COPY=warps"; do
  echo "== $v"; env $v timeout 300 python tools/e2e_probe.py dragons 1920 1080 2>&1 | tail -2
done
RTC_HOST_COPY=dma timeout 300 python tools/e2e_probe.py teapot 1920 1080 2>&1 | tail -2
RTC_HOST_COPY=dma timeout 300 python tools/e2e_probe.py cover 1920 1080 2>&1 | tail -2
