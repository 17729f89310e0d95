#!/bin/bash
cd "$(dirname "$0")/.."
cat > /tmp/dbg3.py <<'PY'
import importlib, sys, time, os
sys.path.insert(0, ".")
import torch
from tests import scenes
pkg = importlib.import_module("ray-tracer-challenge_b200")
text, assets = scenes.shipped_scene("csg")
info = pkg.parse_scene(text, assets, 3840, 2160)
cam, wld = info.camera, info.world
buf = torch.empty((2160, 3840, 3), dtype=torch.float64, device="cuda")
nd = int(sys.argv[1])
for i in range(nd):
    cam.render_device(wld, buf.data_ptr(), 5)
host = pkg.pinned_empty((2160, 3840, 3))
ts = []
for i in range(10):
    st = cam.render(wld, 5, None, out=host).stats
    ts.append(round(st.kernel_ms, 2))
print("device frames before:", nd, "buckets", os.environ.get("RTC_COPY_BUCKETS"), "e2e kernel ms:", ts)
PY
for b in 0 1; do for nd in 0 1 3 13; do RTC_COPY_BUCKETS=$b python /tmp/dbg3.py $nd 2>&1 | tail -1; done; done
