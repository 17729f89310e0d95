#!/bin/bash
# round-2 GPU call 2: work sharing in the mesh traversal -- correctness, variants, probe, ncu
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/run2_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run2_pytest.log
tail -3 gpurun_out/run2_pytest.log
rm -f gpurun_out/variants.jsonl
timeout 900 python tools/kernel_variants.py --only dragons,teapot,nefertiti --out gpurun_out/run2_variants.jsonl \
  steal:: nosteal:variants/nosteal.so: steal2:variants/steal2.so: steal3:variants/steal3.so: 2>&1 | tee gpurun_out/run2_variants.log
for v in "" variants/nosteal.so variants/steal2.so; do
  RTC_B200_LIB=${v:+$PWD/$v} timeout 300 python tools/scaling_probe.py --scene dragons --parts 4,8 2>&1 | tee -a gpurun_out/run2_probe.log
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 5 -c 1 -f -o gpurun_out/r2_b_dragons \
  python tools/bench_all.py --only dragons --reps 6 > gpurun_out/run2_ncu.log 2>&1
echo done
