#!/bin/bash
# 8 GPUs: bench under torchrun at N = 4 and 8 (+ replicated-scene test on 4 devices)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi -L | wc -l
for n in 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/run10_bench$n.json 2> gpurun_out/run10_bench$n.err; echo "bench N=$n rc $?"
  tail -3 gpurun_out/run10_bench$n.err
done
timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "replicated" > gpurun_out/run10_pytest.log 2>&1; tail -3 gpurun_out/run10_pytest.log
python - <<'PY'
import json
for n in (4,8):
    try:
        d=json.loads(open('gpurun_out/run10_bench%d.json'%n).read().strip().splitlines()[-1])
    except Exception as e:
        print(n, "no line", e); continue
    print(n, "value", round(d["value"]), "ms", round(d["ms_per_step"],4), "kernel", [round(x,3) for x in d["kernel_ms_per_rank"]["all"]], "overhead", round(d["sync_overhead_ms"],4), "e2e", round(d["e2e"]["value"]), round(d["e2e"]["ms_per_step"],3))
PY
