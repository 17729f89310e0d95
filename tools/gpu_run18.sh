#!/bin/bash
# round-2 evidence: whole-frame parity report, ncu captures of the kernels that serve the BASELINE configs, launch list
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/parity_r2.jsonl
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/run18_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run18_pytest.log
tail -4 gpurun_out/run18_pytest.log
wc -l gpurun_out/parity_r2.jsonl
NCU="ncu --set full --clock-control none --import-source on -f"
# whole frames: frames 0-2 megakernel, 3-4 wavefront, 5.. the chosen one
timeout 300 $NCU -k regex:render_kernel -s 4 -c 1 -o gpurun_out/r2_c_dragons python tools/ncu_one.py dragons 1920 1080 8 > gpurun_out/run18_ncu_dragons.log 2>&1
timeout 300 $NCU -k regex:render_kernel -s 4 -c 1 -o gpurun_out/r2_c_cover python tools/ncu_one.py cover 1920 1080 8 > gpurun_out/run18_ncu_cover.log 2>&1
timeout 300 $NCU -k regex:render_kernel -s 4 -c 1 -o gpurun_out/r2_c_csg python tools/ncu_one.py csg 3840 2160 8 > gpurun_out/run18_ncu_csg.log 2>&1
timeout 300 $NCU -k regex:render_kernel -s 4 -c 1 -o gpurun_out/r2_c_skybox python tools/ncu_one.py skybox 3840 2160 8 > gpurun_out/run18_ncu_skybox.log 2>&1
timeout 300 $NCU -k regex:render_kernel -s 4 -c 1 -o gpurun_out/r2_c_teapot python tools/ncu_one.py teapot 1920 1080 8 > gpurun_out/run18_ncu_teapot.log 2>&1
# the wavefront kernel on the config it serves: cover, 1/8 row partition (level 1 launch: closest hits of level 1 + shadow queries of level 0)
RTC_KERNEL=wave timeout 300 $NCU -k regex:wave_kernel -s 8 -c 1 -o gpurun_out/r2_c_cover_wave python tools/ncu_one.py cover 1920 1080 4 3 8 > gpurun_out/run18_ncu_wave.log 2>&1
tail -2 gpurun_out/run18_ncu_*.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 5 --warmup 3 --no-configs --no-cpu-baseline > gpurun_out/run18_launches.log 2>&1
ls -la gpurun_out/*.ncu-rep gpurun_out/r2_launches.csv
