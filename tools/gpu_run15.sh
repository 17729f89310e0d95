#!/bin/bash
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/run15_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run15_pytest.log
tail -4 gpurun_out/run15_pytest.log
for sc in dragons cover teapot; do
  python tools/scaling_probe.py --scene $sc --parts 8 --dealt 2>&1 | tee -a gpurun_out/run15_probe.log
  python tools/scaling_probe.py --scene $sc --parts 8 2>&1 | tee -a gpurun_out/run15_probe.log
done
python tools/bench_all.py --reps 5 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except ValueError: continue
    print(d['scene'], d['kernel_ms'], d['Mrays_s'], d['kernel'])
" | tee gpurun_out/run15_bench_all.log
