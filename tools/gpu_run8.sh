#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/run8_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run8_pytest.log
tail -4 gpurun_out/run8_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/run8_bench1.json 2> gpurun_out/run8_bench1.err; echo "bench rc $?"
tail -5 gpurun_out/run8_bench1.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/run8_bench1.json').read().strip().splitlines()[-1])
for k in ("value","ms_per_step","kernel_ms","sync_overhead_ms","first_frame_ms","cold_first_frame_s","gpu_launches"):
    print(k, d.get(k))
print("e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], "pageable", d["e2e"]["pageable_canvas_ms_per_step"], "floor", d["e2e"]["d2h_copy_floor_ms"])
print("rgba8", d.get("e2e_rgba8"))
print("cpu", d.get("cpu_baseline"))
print("roofline", json.dumps(d["roofline"])[:900])
for k,v in (d.get("configs") or {}).items():
    print(k, {kk: v.get(kk) for kk in ("kernel_ms","first_frame_ms","Mrays_s","e2e_ms","kernel","error")}, (v.get("cpu_baseline") or {}).get("value"), (v.get("cpu_baseline") or {}).get("frame_s"))
print("setup", d["scene_setup"])
PY
