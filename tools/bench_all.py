#!/usr/bin/env python
"""Kernel time / Mrays/s of every BASELINE.json config (and optionally all 14 scenes) on one GPU.
Secondary to bench.py (which is the contract line for dragons 1080p): prints one JSON line per config.
usage: python tools/bench_all.py [--all] [--reps 5]"""
import argparse
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import scenes  # noqa: E402

CONFIGS = [("fresnel", 600, 600), ("cover", 1920, 1080), ("teapot", 1920, 1080), ("dragons", 1920, 1080),
           ("csg", 3840, 2160), ("skybox", 3840, 2160)]
NATIVE = [("cubes", 0, 0), ("cylinders", 0, 0), ("reflection_and_refraction", 0, 0), ("xyz", 0, 0), ("groups", 0, 0),
          ("nefertiti", 0, 0), ("align_check", 0, 0), ("earth", 0, 0)]


# algorithmic work per ray.  Analytic scenes: f64 flops (ray transform 56 + primitive test per
# top-level leaf + 160 shading + 28 per pattern level); mesh scenes: BVH node + triangle bytes.
FLOPS_PER_RAY = {"fresnel": 225.0, "cover": 1604.0}
BYTES_PER_RAY = {"teapot": 992.0, "dragons": 1504.0}


def roofline(name, rays_per_s):
    """Fraction of the MEASURED peak (profiles/r1_alu_l2_peaks.json from tools/peaks, MEASURED_PEAKS.json) the
    algorithmic work per ray amounts to; None for configs SURVEY.md gives no per-ray figure for."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    try:
        alu = json.load(open(os.path.join(root, "profiles", "r1_alu_l2_peaks.json")))
    except OSError:
        return None
    if name in FLOPS_PER_RAY:
        a = FLOPS_PER_RAY[name] * rays_per_s / 1e12
        return {"bound": "fp64", "achieved": round(a, 3), "peak": alu["fp64_fma_tflops"], "unit": "TFLOP/s",
                "frac": round(a / alu["fp64_fma_tflops"], 4), "model": "%g f64 flops/ray" % FLOPS_PER_RAY[name]}
    if name in BYTES_PER_RAY:
        a = BYTES_PER_RAY[name] * rays_per_s / 1e9
        hbm = alu["hbm_copy_gbs"]
        try:
            hbm = json.load(open(os.path.join(root, "MEASURED_PEAKS.json")))["hbm_gbs"]
        except (OSError, KeyError):
            pass
        return {"bound": "hbm", "achieved": round(a, 1), "peak": hbm, "unit": "GB/s", "frac": round(a / hbm, 4),
                "frac_of_l2_copy": round(a / alu["l2_copy_gbs"], 4), "model": "%g B/ray" % BYTES_PER_RAY[name]}
    return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--all", action="store_true")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--only", default="")
    ap.add_argument("--kernel", default="", help="pin RTC_KERNEL=mega|wave (default: the library's measured choice)")
    args = ap.parse_args()
    if args.kernel:
        os.environ["RTC_KERNEL"] = args.kernel
    import torch
    pkg = importlib.import_module("ray-tracer-challenge_b200")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    todo = CONFIGS + (NATIVE if args.all else [])
    if args.only:
        todo = [c for c in todo if c[0] in args.only.split(",")]
    for name, w, h in todo:
        text, assets = scenes.shipped_scene(name)
        t0 = time.perf_counter()
        info = pkg.parse_scene(text, assets, w, h)
        setup = time.perf_counter() - t0
        cam = info.camera
        buf = torch.empty((cam.vsize, cam.hsize, 3), dtype=torch.float64, device="cuda")
        times, st = [], None
        for i in range(args.reps + 4):  # 4 untimed frames: the library measures its two kernels on them
            flush.zero_()
            torch.cuda.synchronize()
            st = cam.render_device(info.world, buf.data_ptr())
            if i >= 4:
                times.append(st.kernel_ms)
        times.sort()
        ms = times[len(times) // 2]
        roof = roofline(name, st.rays / (ms * 1e-3))
        print(json.dumps({"roofline": roof, "scene": name, "width": cam.hsize, "height": cam.vsize, "kernel_ms": round(ms, 4),
                          "min_ms": round(times[0], 4), "rays": st.rays, "Mrays_s": round(st.rays / ms / 1e3, 1),
                          "primary": st.primary_rays, "shadow": st.shadow_rays, "secondary": st.secondary_rays, "shadow_culled": st.shadow_rays_culled,
                          "kernel": ("megakernel", "wavefront", "megakernel, several lanes per pixel")[st.kernel] + " (%d launches)" % st.launches, "setup_s": round(setup, 3), "checksum": float(buf.sum().item())}), flush=True)
        info.world.close()


if __name__ == "__main__":
    main()
