#!/bin/bash
# final 1-GPU evidence run of round 2: tests, parity report, bench line, ncu captures + launch list, set-up timing
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/parity_r2.jsonl
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/final_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/final_pytest.log
tail -4 gpurun_out/final_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/final_bench1.json 2> gpurun_out/final_bench1.err; echo "bench rc $?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_ref.json 2> gpurun_out/final_ref.err; echo "ref rc $?"
NCU="ncu --set full --clock-control none --import-source on -f"
timeout 300 $NCU -k regex:render_kernel -s 4 -c 1 -o gpurun_out/r2_d_dragons python tools/ncu_one.py dragons 1920 1080 8 > gpurun_out/final_ncu_dragons.log 2>&1
timeout 300 $NCU -k regex:render_kernel -s 4 -c 1 -o gpurun_out/r2_d_teapot python tools/ncu_one.py teapot 1920 1080 8 > gpurun_out/final_ncu_teapot.log 2>&1
timeout 300 $NCU -k regex:render_kernel -s 4 -c 1 -o gpurun_out/r2_d_fresnel python tools/ncu_one.py fresnel 600 600 8 > gpurun_out/final_ncu_fresnel.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 5 --warmup 3 --no-configs --no-cpu-baseline > gpurun_out/final_launches.log 2>&1
for m in host gpu; do for sc in nefertiti dragons; do RTC_BVH=$m RTC_LOAD_TIMING=1 python tools/ncu_one.py $sc 640 360 1 2>&1 | grep -E "scene build|load:" | sed "s/^/$m $sc /"; done; done > gpurun_out/final_setup_timing.log 2>&1
cat gpurun_out/final_setup_timing.log | grep "BVH build"
python tools/bench_all.py --all --reps 5 > gpurun_out/r2_bench_all.jsonl 2>gpurun_out/final_bench_all.err
python tools/scaling_probe.py --scene dragons --parts 1,2,4,8 > gpurun_out/r2_scaling_probe_rows.jsonl 2>/dev/null
python tools/scaling_probe.py --scene dragons --parts 2,4,8 --dealt > gpurun_out/r2_scaling_probe_dealt.jsonl 2>/dev/null
ls -la gpurun_out/final_bench1.json gpurun_out/r2_d_*.ncu-rep
