#!/bin/bash
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests -m gpu -x -q -k "shared_tile or lanes_per_pixel or small_partitions" > gpurun_out/run13_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run13_pytest.log
tail -3 gpurun_out/run13_pytest.log
for sh in 0 3 5; do
for sc in dragons cover teapot; do
  echo "shift $sh"; RTC_DEAL_SHIFT=$sh python tools/scaling_probe.py --scene $sc --parts 8 --dealt 2>&1 | tee -a gpurun_out/run13_probe.log
done; done
for sc in cover teapot; do python tools/scaling_probe.py --scene $sc --parts 8 2>&1 | tee -a gpurun_out/run13_probe.log; done
