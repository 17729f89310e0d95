#!/bin/bash
cd "$(dirname "$0")/.."
RTC_TRACE_FRAMES=1 python tools/scaling_probe.py --scene cover --parts 8 --dealt 2>&1 | awk 'NR<=40 || /scene/' | tee gpurun_out/run16_trace_dealt.log
RTC_TRACE_FRAMES=1 python tools/scaling_probe.py --scene cover --parts 8 2>&1 | awk 'NR<=16 || /scene/' | tee gpurun_out/run16_trace_rows.log
