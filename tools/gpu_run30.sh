#!/bin/bash
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/run30_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run30_pytest.log
tail -3 gpurun_out/run30_pytest.log
timeout 600 python tools/kernel_variants.py --only dragons,teapot,fresnel,skybox --out gpurun_out/run30_variants.jsonl --reps 7 box:: nobox::RTC_SCENE_BOX=0 2>&1 | tee gpurun_out/run30_variants.log
