#!/bin/bash
# bench.py under torchrun at N = $1 GPUs -> gpurun_out/r2_scale_${N}gpu.json
cd "$(dirname "$0")/.."
n=$1
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2953$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r2_scale_${n}gpu.json 2> gpurun_out/r2_scale_${n}gpu.err; echo "bench N=$n rc $?"
python - <<PY
import json
d=json.loads(open('gpurun_out/r2_scale_${n}gpu.json').read().strip().splitlines()[-1])
print($n, "value", round(d["value"]), "ms", round(d["ms_per_step"],4), "kernel", [round(x,3) for x in d["kernel_ms_per_rank"]["all"]], "overhead", round(d["sync_overhead_ms"],4), "e2e", round(d["e2e"]["value"]), round(d["e2e"]["ms_per_step"],3))
PY
