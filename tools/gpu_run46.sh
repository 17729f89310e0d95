#!/bin/bash
# 2 GPUs: GPU tests that need two devices + bench at N = 2 (e2e leg: row partitions into the shared pinned canvas through the copy engine)
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests -m gpu -x -q -k "replicated or shared or world or multi" > gpurun_out/run46_pytest.log 2>&1; tail -3 gpurun_out/run46_pytest.log
for mode in dma warps; do
RTC_HOST_COPY=$mode timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/run46_bench2_$mode.json 2> gpurun_out/run46_bench2_$mode.err; echo "bench N=2 $mode rc $?"
python - <<PY
import json
d=json.loads(open('gpurun_out/run46_bench2_$mode.json').read().strip().splitlines()[-1])
print("value", round(d["value"]), "ms", round(d["ms_per_step"],4), "kernel", [round(x,3) for x in d["kernel_ms_per_rank"]["all"]], "e2e", round(d["e2e"]["value"]), round(d["e2e"]["ms_per_step"],3))
PY
done
