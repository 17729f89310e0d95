"""Debug helper (run on the GPU box): dumps divergent pixels of a scene."""
import importlib, sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests import scenes
from oracle import oracle_py
pkg = importlib.import_module("ray-tracer-challenge_b200")

def run(label, text, asset_dir, w, h):
    info = pkg.parse_scene(text, asset_dir, w, h)
    c = info.camera.render(info.world)
    ref = oracle_py.OracleScene(json_text=text, asset_dir=asset_dir, width=w, height=h)
    want, counts = ref.render()
    diff = np.abs(c.pixels - want).max(axis=2)
    bad = np.argwhere(diff > 1e-4)
    st = c.stats
    print(label, w, h, "divergent", len(bad), "rays", (st.primary_rays, st.shadow_rays, st.secondary_rays), counts, "ms", st.kernel_ms)
    for y, x in bad[:40]:
        print("  px", x, y, "gpu", c.pixels[y, x], "ref", want[y, x])
    np.save("gpurun_out/%s_gpu.npy" % label, c.pixels)
    np.save("gpurun_out/%s_ref.npy" % label, want)

if __name__ == "__main__":
    import tempfile
    d = tempfile.mkdtemp()
    scenes.write_mesh_obj(os.path.join(d, "torus.obj"), smooth=True, groups=False)
    run("mesh_smooth", scenes.mesh_scene("torus.obj"), d, 160, 100)
    for name in sys.argv[1:]:
        s = scenes.shipped_scene(name)
        import json
        cam = json.loads(s[0])["camera"]
        run(name, s[0], s[1], cam["width"] // 4, cam["height"] // 4)
