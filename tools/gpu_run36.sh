#!/bin/bash
# f32 triangle pre-test: GPU tests (bit-identity, parity), kernel times of the BASELINE configs, dealt shares
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/run36_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run36_pytest.log
tail -4 gpurun_out/run36_pytest.log
timeout 600 python tools/bench_all.py --reps 7 2>&1 | tee gpurun_out/run36_bench_all.jsonl | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except ValueError: continue
    print(d['scene'], d.get('kernel'), round(d['kernel_ms'], 4), round(d['Mrays_s']))
"
for sc in dragons teapot; do timeout 300 python tools/scaling_probe.py --dealt --scene $sc --parts 2,4,8 2>&1 | tail -3; done | tee gpurun_out/run36_dealt.jsonl
