#!/bin/bash
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests -m gpu -x -q -k "shared or replicated or async or static" 2>&1 | tail -2
for v in "" "RTC_FINE_ORDER=0"; do echo "== ${v:-fine order}"; for sc in dragons teapot cover; do env $v timeout 300 python tools/scaling_probe.py --dealt --scene $sc --parts 2,4,8 2>/dev/null | cut -c1-180; done; done
