#!/bin/bash
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/run12_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run12_pytest.log
tail -4 gpurun_out/run12_pytest.log
python tools/scaling_probe.py --scene dragons --parts 2,4,8 --dealt 2>&1 | tee gpurun_out/run12_probe.log
python tools/scaling_probe.py --scene cover --parts 8 --dealt 2>&1 | tee -a gpurun_out/run12_probe.log
python tools/scaling_probe.py --scene teapot --parts 8 --dealt 2>&1 | tee -a gpurun_out/run12_probe.log
