#!/bin/bash
cd "$(dirname "$0")/.."
RTC_TRACE_FRAMES=1 python - <<'PY' 2>&1 | tail -30
import importlib, sys, time
sys.path.insert(0, ".")
import torch
from tests import scenes
pkg = importlib.import_module("ray-tracer-challenge_b200")
text, assets = scenes.shipped_scene("csg")
info = pkg.parse_scene(text, assets, 3840, 2160)
cam, wld = info.camera, info.world
buf = torch.empty((2160, 3840, 3), dtype=torch.float64, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for i in range(13):
    flush.zero_(); torch.cuda.synchronize()
    cam.render_device(wld, buf.data_ptr(), 5)
print("--- e2e", file=sys.stderr)
host = pkg.pinned_empty((2160, 3840, 3))
for i in range(8):
    t = time.perf_counter()
    cam.render(wld, 5, None, out=host)
    print("e2e frame %d %.3f ms" % (i, (time.perf_counter() - t) * 1e3), file=sys.stderr)
PY
