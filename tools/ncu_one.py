#!/usr/bin/env python
"""One scene rendered N times into device memory -- the command the ncu captures of profiles/ wrap.
usage: python tools/ncu_one.py scene width height [frames] [partition part n_parts]"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import scenes  # noqa: E402


def main():
    import torch
    name, w, h = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    frames = int(sys.argv[4]) if len(sys.argv) > 4 else 8
    part = (8, int(sys.argv[5]), int(sys.argv[6])) if len(sys.argv) > 6 else None
    pkg = importlib.import_module("ray-tracer-challenge_b200")
    text, assets = scenes.shipped_scene(name)
    info = pkg.parse_scene(text, assets, w, h)
    cam = info.camera
    buf = torch.empty((cam.vsize, cam.hsize, 3), dtype=torch.float64, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    st = None
    for _ in range(frames):
        flush.zero_()
        torch.cuda.synchronize()
        st = cam.render_device_frame(info.world, buf.data_ptr(), 5, part)
    print(name, cam.hsize, cam.vsize, "kernel_ms", st.kernel_ms, "kernel", st.kernel, "rays", st.rays)


if __name__ == "__main__":
    main()
