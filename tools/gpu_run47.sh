#!/bin/bash
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/run47_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run47_pytest.log
tail -4 gpurun_out/run47_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/run47_bench.json 2> gpurun_out/run47_bench.err; echo "bench rc $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/run47_bench.json').read().strip().splitlines()[-1])
print("value", round(d["value"]), "ms", round(d["ms_per_step"],4), "kernel", d.get("kernel_ms"), "e2e", round(d["e2e"]["value"]), d["e2e"].get("ms_per_step"), "rgba8", d["e2e_rgba8"]["ms_per_step"])
for k,v in d.get("configs",{}).items():
    print(k, {a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a in ("kernel_ms","value","e2e_ms")})
PY
