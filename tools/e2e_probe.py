#!/usr/bin/env python
"""End-to-end frame time into a pinned host canvas (rtc_render), decoupled copy warps against direct stores.
usage: python tools/e2e_probe.py [scene w h]   (RTC_COPY_CTAS=0 -> the direct-store path)"""
import importlib
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import scenes  # noqa: E402


def main():
    import numpy as np
    import torch
    name = sys.argv[1] if len(sys.argv) > 1 else "dragons"
    w, h = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1920, 1080)
    pkg = importlib.import_module("ray-tracer-challenge_b200")
    text, assets = scenes.shipped_scene(name)
    info = pkg.parse_scene(text, assets, w, h)
    cam, wld = info.camera, info.world
    host = pkg.pinned_empty((h, w, 3))
    ref = cam.render(wld).pixels  # pageable: staged
    for _ in range(8):
        st = cam.render(wld, out=host).stats
    assert np.array_equal(host, ref), "pinned canvas differs from the staged frame"
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = 30
    k_ms = 0.0
    for _ in range(n):
        st = cam.render(wld, out=host).stats
        k_ms += st.kernel_ms
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / n * 1e3
    host[:] = 0.0  # the steady-state path really delivers the frame
    st = cam.render(wld, out=host).stats
    assert np.array_equal(host, ref), "steady-state frame did not reach the pinned canvas"
    print("%s %dx%d host_copy=%s copy_ctas=%s: e2e %.4f ms/frame, kernel %.4f ms (%s), %.1f Mrays/s" % (
        name, w, h, os.environ.get("RTC_HOST_COPY", "default"), os.environ.get("RTC_COPY_CTAS", "default"), ms, k_ms / n,
        ("mega", "wave", "mixed")[st.kernel], st.rays / ms / 1e3))
    # the lib.zig twin: RGBA8 into pinned memory
    import ctypes
    staged, _ = cam.render_rgba8(wld)
    pin8 = pkg.pinned_empty((h, w, 4), dtype=np.uint8)
    stt = pkg.rtc_stats()

    def once():
        pkg._check(pkg.lib().rtc_render_rgba8(wld._h, ctypes.byref(cam.c), 5, None, pin8.ctypes.data, ctypes.byref(stt)))
    for _ in range(8):
        once()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        once()
    ms8 = (time.perf_counter() - t0) / n * 1e3
    pin8[:] = 0
    once()
    assert np.array_equal(pin8, staged), "RGBA8 frame in pinned memory differs from the staged one"
    print("%s %dx%d rgba8: e2e %.4f ms/frame, kernel %.4f ms" % (name, w, h, ms8, stt.kernel_ms))


if __name__ == "__main__":
    main()
