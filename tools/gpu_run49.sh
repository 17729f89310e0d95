#!/bin/bash
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests -m gpu -x -q -k "shared or lpp or lanes or async or static or kernels" 2>&1 | tail -2
for sc in dragons teapot cover; do RTC_TRACE_FRAMES=1 timeout 300 python tools/scaling_probe.py --dealt --scene $sc --parts 2,4,8 2> gpurun_out/run49_$sc.err | cut -c1-200; done
grep "px 259200 candidate" gpurun_out/run49_dragons.err | awk '{print $7, $NF-1, $(NF-1)}' | sort | uniq -c | sort -k2n | head -40
