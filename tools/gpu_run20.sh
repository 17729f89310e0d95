#!/bin/bash
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/run20_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run20_pytest.log
tail -4 gpurun_out/run20_pytest.log
python tools/bench_all.py --all --reps 5 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except ValueError: continue
    print(d['scene'], d['width'], d['height'], d['kernel_ms'], d['Mrays_s'], d['kernel'])
" | tee gpurun_out/run20_bench_all.log
