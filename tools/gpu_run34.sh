#!/bin/bash
# rtc_render_device_static: its test alone in a fresh process, then the whole GPU suite (it runs last), then a bench line
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/run34_alone.log 2>&1; echo "alone exit $?" >> gpurun_out/run34_alone.log
tail -15 gpurun_out/run34_alone.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/run34_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/run34_pytest.log
tail -5 gpurun_out/run34_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/run34_bench.json 2> gpurun_out/run34_bench.err; echo "bench rc $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/run34_bench.json').read().strip().splitlines()[-1])
print("value", round(d["value"]), "ms", round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"]), d["e2e"].get("ms_per_step"))
PY
