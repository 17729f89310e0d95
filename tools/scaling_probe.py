#!/usr/bin/env python
"""Predicts the multi-GPU kernel scaling on ONE GPU: renders every partition (tile_rows, r, N) of the bench
frame separately (L2 flushed before each) and prints max/mean kernel time over the parts for N = 1, 2, 4, 8.
usage: python tools/scaling_probe.py [--scene dragons] [--tile-rows 8]"""
import argparse
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import scenes  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="dragons")
    ap.add_argument("--tile-rows", type=int, default=8)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--parts", default="1,2,4,8")
    ap.add_argument("--kernel", default="", help="pin RTC_KERNEL=mega|wave")
    a = ap.parse_args()
    if a.kernel:
        os.environ["RTC_KERNEL"] = a.kernel
    import torch
    pkg = importlib.import_module("ray-tracer-challenge_b200")
    text, assets = scenes.shipped_scene(a.scene)
    info = pkg.parse_scene(text, assets, a.width, a.height)
    cam = info.camera
    frame = torch.empty((cam.vsize, cam.hsize, 3), dtype=torch.float64, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for n in [int(x) for x in a.parts.split(",")]:
        per = []
        for r in range(min(n, 8)):
            best = 1e9
            for _ in range(10):
                if not a.no_flush:
                    flush.zero_()
                torch.cuda.synchronize()
                st = cam.render_device_frame(info.world, frame.data_ptr(), 5, (a.tile_rows, r, n))
                best = min(best, st.kernel_ms)
            per.append(best)
        print(json.dumps({"scene": a.scene, "n_parts": n, "tile_rows": a.tile_rows, "max_ms": round(max(per), 4),
                          "mean_ms": round(sum(per) / len(per), 4), "speedup_vs_1": None, "parts": [round(p, 3) for p in per]}),
              flush=True)


if __name__ == "__main__":
    main()
