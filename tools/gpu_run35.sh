#!/bin/bash
# ray-tree splitting on heavy tiles (-DRTC_HEAVY_GROUPS=2|4|8): identity check + dealt 1/4 and 1/8 shares on one GPU
cd "$(dirname "$0")/.."
for v in "" g2 g4 g8; do
  if [ -n "$v" ]; then export RTC_B200_LIB=$PWD/variants/librtc_$v.so; else unset RTC_B200_LIB; fi
  echo "== variant ${v:-default}"
  timeout 300 python tools/lpp_check.py dragons 2>&1 | grep '"lpp": 0' 
  for sc in dragons teapot cover; do
    timeout 300 python tools/scaling_probe.py --dealt --scene $sc --parts 4,8 2>&1 | tail -2
  done
done 2>&1 | tee gpurun_out/run35.log
