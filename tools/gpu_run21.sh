#!/bin/bash
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests -m gpu -x -q -k "lanes_per_pixel or shared_tile or small_partitions" > gpurun_out/run21_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run21_pytest.log
tail -4 gpurun_out/run21_pytest.log
grep -q "pytest exit 0" gpurun_out/run21_pytest.log || exit 1
for v in "" variants/heavy4.so; do
  echo "lib=$v" | tee -a gpurun_out/run21_probe.log
  RTC_B200_LIB=${v:+$PWD/$v} python tools/scaling_probe.py --scene dragons --parts 4,8 --dealt 2>&1 | tee -a gpurun_out/run21_probe.log
  RTC_B200_LIB=${v:+$PWD/$v} RTC_LPT_ALWAYS=1 RTC_LPP=0 python tools/scaling_probe.py --scene dragons --parts 8 --dealt 2>&1 | tee -a gpurun_out/run21_probe.log
done
