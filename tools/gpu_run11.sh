#!/bin/bash
cd "$(dirname "$0")/.."
RTC_LPT_ALWAYS=1 python tools/scaling_probe.py --scene dragons --parts 8 --dealt 2>&1 | tee gpurun_out/run11_probe.log
RTC_LPT_ALWAYS=1 RTC_LPP=1 python tools/scaling_probe.py --scene dragons --parts 8 --dealt 2>&1 | tee -a gpurun_out/run11_probe.log
RTC_LPT_ALWAYS=1 RTC_LPP=4 python tools/scaling_probe.py --scene dragons --parts 8 --dealt 2>&1 | tee -a gpurun_out/run11_probe.log
