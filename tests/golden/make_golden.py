"""Generates tests/golden/*.npz: small f64 renders by the CPU oracle (oracle/, the C++ restatement of
the reference's algorithm) of every shipped scene and of the synthetic test scenes, with ray counts.

Run where /root/reference is mounted (assets are staged by __graft_entry__.build()):
    python tests/golden/make_golden.py
The reference itself is Zig and cannot be executed here, so these vectors pin the ORACLE's output
(itself pinned to the reference by oracle/kat_tests.cpp); they are regression fixtures for the
oracle and an oracle-free check for the CUDA path on the GPU box."""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import oracle_py  # noqa: E402
from tests import scenes  # noqa: E402

SIZES = {"align_check": (80, 40), "cover": (64, 64), "csg": (80, 45), "cubes": (80, 40), "cylinders": (80, 40),
         "dragons": (75, 30), "earth": (80, 40), "fresnel": (60, 60), "groups": (90, 30), "nefertiti": (36, 60),
         "reflection_and_refraction": (80, 40), "skybox": (80, 40), "teapot": (75, 45), "xyz": (80, 45)}


def render(text, assets, w, h):
    s = oracle_py.OracleScene(json_text=text, asset_dir=assets, width=w, height=h)
    img, c = s.render()
    return img, np.array([c["primary"], c["shadow"], c["secondary"]], dtype=np.int64)


def main():
    out = {}
    for name, (w, h) in SIZES.items():
        text, assets = scenes.shipped_scene(name)
        img, c = render(text, assets, w, h)
        out[name] = img
        out[name + "__rays"] = c
        print(name, w, h, c, float(img.mean()))
    img, c = render(scenes.SMOKE_SCENE, "", 96, 64)
    out["smoke"], out["smoke__rays"] = img, c
    d = tempfile.mkdtemp()
    scenes.write_mesh_obj(os.path.join(d, "torus.obj"), smooth=True, groups=False)
    img, c = render(scenes.mesh_scene("torus.obj"), d, 90, 60)
    out["mesh_smooth"], out["mesh_smooth__rays"] = img, c
    np.savez_compressed(os.path.join(HERE, "oracle_renders.npz"), **out)


if __name__ == "__main__":
    main()
