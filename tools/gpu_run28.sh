#!/bin/bash
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests -m gpu -x -q -k "gpu_built" > gpurun_out/run28_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run28_pytest.log
tail -12 gpurun_out/run28_pytest.log
for m in host gpu; do RTC_BVH=$m python tools/bench_all.py --all --only teapot,dragons,nefertiti --reps 5 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except ValueError: continue
    print('$m', d['scene'], d['width'], d['height'], d['kernel_ms'], d['Mrays_s'], 'setup_s', d['setup_s'])
"; done | tee gpurun_out/run28_bvh.log
