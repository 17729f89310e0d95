#!/bin/bash
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/run45_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run45_pytest.log
tail -4 gpurun_out/run45_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/run45_bench.json 2> gpurun_out/run45_bench.err; echo "bench rc $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/run45_bench.json').read().strip().splitlines()[-1])
print("value", round(d["value"]), "ms", round(d["ms_per_step"],4), "kernel", d.get("kernel_ms"), "e2e", round(d["e2e"]["value"]), d["e2e"].get("ms_per_step"), "rgba8", d.get("e2e_rgba8"))
for k,v in d.get("configs",{}).items():
    print(k, {a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a in ("kernel_ms","value","e2e_ms","e2e_value","e2e")})
PY
RTC_HOST_COPY=dma timeout 300 python tools/e2e_probe.py csg 3840 2160 2>&1 | tail -2
