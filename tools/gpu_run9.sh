#!/bin/bash
# 2 GPUs: replicated-scene test (rtc_set_devices) + bench under torchrun
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "replicated or shared_tile" > gpurun_out/run9_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run9_pytest.log
tail -15 gpurun_out/run9_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/run9_bench2.json 2> gpurun_out/run9_bench2.err; echo "bench rc $?"
tail -8 gpurun_out/run9_bench2.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/run9_bench2.json').read().strip().splitlines()[-1])
for k in ("value","ms_per_step","kernel_ms","kernel_ms_per_rank","sync_overhead_ms","gpu_launches","exchange"):
    print(k, d.get(k))
print("e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"])
print("setup", d["scene_setup"])
PY
