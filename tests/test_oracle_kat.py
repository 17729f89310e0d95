"""Pins the oracle: (1) the reference's own known-answer tests, transcribed in oracle/kat_tests.cpp;
(2) the committed golden renders (tests/golden/oracle_renders.npz, made by make_golden.py)."""
import os
import subprocess
import tempfile

import numpy as np
import pytest

from tests import scenes
from tests.golden import make_golden

GOLDEN = os.path.join(scenes.ROOT, "tests", "golden", "oracle_renders.npz")


def test_reference_known_answer_tests(oracle):
    exe = os.path.join(scenes.ROOT, "oracle", "kat_tests")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", os.path.join(scenes.ROOT, "oracle"), "kat_tests"])
    out = subprocess.run([exe], capture_output=True, text=True)
    print(out.stdout)
    assert out.returncode == 0, out.stdout
    assert "0 failures" in out.stdout


def test_oracle_matches_golden_synthetic(oracle):
    g = np.load(GOLDEN)
    s = oracle.OracleScene(json_text=scenes.SMOKE_SCENE, width=96, height=64)
    img, c = s.render()
    assert np.abs(img - g["smoke"]).max() <= 1e-12
    assert [c["primary"], c["shadow"], c["secondary"]] == list(g["smoke__rays"])
    d = tempfile.mkdtemp()
    scenes.write_mesh_obj(os.path.join(d, "torus.obj"), smooth=True, groups=False)
    s = oracle.OracleScene(json_text=scenes.mesh_scene("torus.obj"), asset_dir=d, width=90, height=60)
    img, c = s.render()
    assert np.abs(img - g["mesh_smooth"]).max() <= 1e-12


@pytest.mark.parametrize("name", ["fresnel", "csg", "cover", "groups", "earth", "teapot"])
def test_oracle_matches_golden_shipped(oracle, name):
    s = scenes.shipped_scene(name)
    if s is None:
        pytest.skip("reference assets not staged (assets/_ref)")
    g = np.load(GOLDEN)
    w, h = make_golden.SIZES[name]
    sc = oracle.OracleScene(json_text=s[0], asset_dir=s[1], width=w, height=h)
    img, c = sc.render()
    assert np.abs(img - g[name]).max() <= 1e-12
    assert [c["primary"], c["shadow"], c["secondary"]] == list(g[name + "__rays"])
