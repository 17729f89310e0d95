#!/bin/bash
cd "$(dirname "$0")/.."
RTC_TRACE_FRAMES=1 python tools/scaling_probe.py --scene teapot --parts 8 2>&1 | head -60 | tee gpurun_out/run14_trace.log
