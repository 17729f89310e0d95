#!/bin/bash
# 4 GPUs: partition frames through the copy warps (e2e leg), bench at N = 2 and 4; GPU tests first (1 GPU subset runs on device 0)
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/run31_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run31_pytest.log
tail -3 gpurun_out/run31_pytest.log
for n in 2 4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2954$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/run31_bench$n.json 2> gpurun_out/run31_bench$n.err; echo "bench N=$n rc $?"
  tail -2 gpurun_out/run31_bench$n.err
done
python - <<'PY'
import json
for n in (2,4):
    try:
        d=json.loads(open('gpurun_out/run31_bench%d.json'%n).read().strip().splitlines()[-1])
    except Exception as e:
        print(n, "no line", e); continue
    print(n, "value", round(d["value"]), "ms", round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"]), round(d["e2e"]["ms_per_step"],3))
PY
