#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export RTC_KERNEL=mega
timeout 600 compute-sanitizer --tool memcheck --print-limit 5 python -m pytest tests/test_parity_gpu.py -k "mesh_scene" -x -q > gpurun_out/dbg_memcheck.log 2>&1
tail -60 gpurun_out/dbg_memcheck.log
