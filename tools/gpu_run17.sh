#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for n in 8 4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/run17_bench$n.json 2> gpurun_out/run17_bench$n.err; echo "bench N=$n rc $?"
done
python - <<'PY'
import json
for n in (4,8):
    try:
        d=json.loads(open('gpurun_out/run17_bench%d.json'%n).read().strip().splitlines()[-1])
    except Exception as e:
        print(n, "no line", e); continue
    print(n, "value", round(d["value"]), "ms", round(d["ms_per_step"],4), "kernel", [round(x,3) for x in d["kernel_ms_per_rank"]["all"]], "overhead", round(d["sync_overhead_ms"],4), "e2e", round(d["e2e"]["value"]), round(d["e2e"]["ms_per_step"],3))
PY
