#!/bin/bash
# final multi-GPU evidence run of round 2: bench.py under torchrun at N = 2, 4, 8 (+ reference arm smoke, replica test)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for n in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2953$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r2_scale_${n}gpu.json 2> gpurun_out/r2_scale_${n}gpu.err; echo "bench N=$n rc $?"
done
timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "replicated" > gpurun_out/final8_pytest.log 2>&1; tail -2 gpurun_out/final8_pytest.log
python - <<'PY'
import json
for n in (2,4,8):
    try:
        d=json.loads(open('gpurun_out/r2_scale_%dgpu.json'%n).read().strip().splitlines()[-1])
    except Exception as e:
        print(n, "no line", e); continue
    print(n, "value", round(d["value"]), "ms", round(d["ms_per_step"],4), "kernel", [round(x,3) for x in d["kernel_ms_per_rank"]["all"]], "overhead", round(d["sync_overhead_ms"],4), "e2e", round(d["e2e"]["value"]), round(d["e2e"]["ms_per_step"],3), "setup", {k: round(v,3) for k,v in d["scene_setup"].items() if isinstance(v,float)})
PY
