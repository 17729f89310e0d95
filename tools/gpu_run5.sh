#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/run5_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run5_pytest.log
tail -5 gpurun_out/run5_pytest.log
grep -q "pytest exit 0" gpurun_out/run5_pytest.log || exit 1
python tools/scaling_probe.py --scene dragons --parts 1,2,4,8 2>&1 | tee gpurun_out/run5_probe_rows.log
python tools/scaling_probe.py --scene dragons --parts 2,4,8 --dealt 2>&1 | tee gpurun_out/run5_probe_dealt.log
python tools/bench_all.py --reps 5 2>&1 | tee gpurun_out/run5_bench_all.jsonl
