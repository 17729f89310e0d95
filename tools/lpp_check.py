#!/usr/bin/env python
"""Lanes-per-pixel variants of the megakernel (render_kernel<MESH, LPP>): frames and ray counts must be
bit-identical to LPP=1; prints kernel ms for the whole frame and for a 1/8 partition.
usage: python tools/lpp_check.py [scene ...]"""
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import scenes  # noqa: E402


def main():
    import torch
    pkg = importlib.import_module("ray-tracer-challenge_b200")
    os.environ["RTC_KERNEL"] = "mega"
    for name in sys.argv[1:] or ["dragons", "cover", "fresnel", "teapot", "csg", "skybox"]:
        text, assets = scenes.shipped_scene(name)
        w, h = (600, 600) if name == "fresnel" else (1920, 1080)
        info = pkg.parse_scene(text, assets, w, h)
        cam = info.camera
        frames = {}
        for lpp in (1, 2, 4, 0):  # 0 = mixed: four lanes per pixel on the tiles the last frame found expensive
            os.environ["RTC_LPP"] = str(lpp)
            frame = torch.zeros((cam.vsize, cam.hsize, 3), dtype=torch.float64, device="cuda")
            out = {"scene": name, "lpp": lpp}
            for part in ((8, 0, 1), (8, 3, 8)):
                best = 1e9
                for _ in range(6):
                    st = cam.render_device_frame(info.world, frame.data_ptr(), 5, part)
                    best = min(best, st.kernel_ms)
                out["ms_1_of_%d" % part[2]] = round(best, 4)
                if part[2] == 1:
                    out["rays"] = (st.primary_rays, st.shadow_rays, st.secondary_rays)
            frames[lpp] = (frame.clone(), out["rays"])
            out["identical_to_lpp1"] = bool(torch.equal(frames[1][0], frame)) and frames[1][1] == out["rays"]
            print(json.dumps(out), flush=True)
            assert out["identical_to_lpp1"], "LPP=%d differs" % lpp


if __name__ == "__main__":
    main()
