"""Parity of the CUDA path (through the C ABI) against the CPU oracle, same inputs.

Bar (BASELINE.json north_star): linear RGB within 1e-4 absolute on >= 99.9 % of pixels, quantised
8-bit output within +-1 LSB on those pixels, divergent (epsilon-flip) pixel count reported; ray
counts P/S/Q are deterministic and must match the oracle's up to the divergent pixels."""
import json
import os

import numpy as np
import pytest

from tests import scenes

pytestmark = pytest.mark.gpu
TOL = 1e-4


def compare(pkg, oracle, text, asset_dir, w, h, min_frac=0.999, label=""):
    info = pkg.parse_scene(text, asset_dir, w, h)
    canvas = info.camera.render(info.world)
    ref = oracle.OracleScene(json_text=text, asset_dir=asset_dir, width=w, height=h)
    want, counts = ref.render()
    diff = np.abs(canvas.pixels - want).max(axis=2)
    ok = diff <= TOL
    frac = float(ok.mean())
    q_gpu, q_ref = canvas.to_image().astype(int), oracle.quantise(want).astype(int)
    lsb = np.abs(q_gpu - q_ref).max(axis=2)
    st = canvas.stats
    print("%s %dx%d: within %.0e: %.5f  divergent px: %d  max|d| on ok px: %.2e  8-bit max diff on ok px: %d  "
          "rays gpu (%d,%d,%d) oracle (%d,%d,%d)  kernel %.2f ms" % (
              label, w, h, TOL, frac, int((~ok).sum()), float(diff[ok].max()) if ok.any() else 0.0,
              int(lsb[ok].max()) if ok.any() else 0, st.primary_rays, st.shadow_rays, st.secondary_rays,
              counts["primary"], counts["shadow"], counts["secondary"], st.kernel_ms))
    assert np.isfinite(canvas.pixels).all()
    assert frac >= min_frac
    assert lsb[ok].max() <= 1
    assert st.primary_rays == counts["primary"] == w * h
    # shadow / secondary counts may differ only through divergent pixels (a ray tree has at most 63 nodes): none diverges
    # -> the counts are EQUAL
    n_div = int((~ok).sum())
    assert abs(int(st.shadow_rays) - counts["shadow"]) <= 64 * n_div * 4
    assert abs(int(st.secondary_rays) - counts["secondary"]) <= 64 * n_div
    info.world.close()
    return canvas, want


def test_default_world_kat(pkg, oracle):
    # camera.zig:171-187: pixel (5,5) of the 11x11 default world
    info = pkg.parse_scene(scenes.default_world(), "")
    c = info.camera.render(info.world)
    assert np.allclose(c.pixels[5, 5], [0.38066, 0.47583, 0.2855], atol=1e-5)
    compare(pkg, oracle, scenes.default_world(), "", 11, 11, label="default")


def test_smoke_scene(pkg, oracle):
    compare(pkg, oracle, scenes.SMOKE_SCENE, "", 192, 128, label="smoke")


@pytest.mark.parametrize("smooth,groups", [(True, False), (False, True)])
def test_mesh_scene(pkg, oracle, tmp_path, smooth, groups):
    scenes.write_mesh_obj(str(tmp_path / "torus.obj"), smooth=smooth, groups=groups)
    compare(pkg, oracle, scenes.mesh_scene("torus.obj"), str(tmp_path), 160, 100, label="mesh")


@pytest.mark.parametrize("name,w,h", [
    ("fresnel", 150, 150), ("cover", 160, 160), ("teapot", 125, 75), ("dragons", 125, 50), ("csg", 160, 90),
    ("skybox", 200, 100), ("earth", 200, 100), ("cubes", 150, 75), ("cylinders", 200, 100),
    ("reflection_and_refraction", 200, 100), ("xyz", 150, 75), ("groups", 150, 50), ("nefertiti", 60, 100),
    ("align_check", 200, 100)])
def test_shipped_scene_small(pkg, oracle, name, w, h):
    s = scenes.shipped_scene(name)
    if s is None:
        pytest.skip("reference assets not staged (assets/_ref)")
    compare(pkg, oracle, s[0], s[1], w, h, label=name)


def test_partition_and_rgba8(pkg, oracle):
    text = scenes.SMOKE_SCENE
    info = pkg.parse_scene(text, "", 96, 64)
    full = info.camera.render(info.world).pixels
    acc = np.zeros_like(full)
    rows = 0
    for part in range(3):
        c = info.camera.render(info.world, partition=(8, part, 3))
        rows += info.camera.partition_rows((8, part, 3))
        acc += c.pixels
    assert rows == 64
    assert np.array_equal(acc, full)  # same kernel, same pixels: bit-identical
    rgba, st = info.camera.render_rgba8(info.world)
    assert np.array_equal(rgba[..., :3], pkg.Canvas(full).to_image())
    assert (rgba[..., 3] == 255).all()
    # re-render after a camera move (lib.zig:166-190) matches the oracle too
    info.camera.look_at((3, 2.5, -6), (0, 0.8, 0), (0, 1, 0))
    moved = info.camera.render(info.world).pixels
    ref = oracle.OracleScene(json_text=text, width=96, height=64)
    ref.set_camera((3, 2.5, -6), (0, 0.8, 0), (0, 1, 0))
    want, _ = ref.render()
    assert (np.abs(moved - want).max(axis=2) <= TOL).mean() >= 0.999


def test_render_device_frame_partitions(pkg):
    """Multi-GPU output path on one GPU: every partition rendered straight into ONE full frame (what each rank
    does into rank 0's peer-mapped canvas) reproduces the whole-frame render bit for bit."""
    import torch
    info = pkg.parse_scene(scenes.SMOKE_SCENE, "", 96, 61)  # 61 rows: last tile is ragged
    full = info.camera.render(info.world)
    frame = torch.full((61, 96, 3), float("nan"), dtype=torch.float64, device="cuda")
    rays = 0
    for part in range(4):
        st = info.camera.render_device_frame(info.world, frame.data_ptr(), partition=(8, part, 4))
        rays += st.rays
    torch.cuda.synchronize()
    assert np.array_equal(frame.cpu().numpy(), full.pixels)
    assert rays == full.stats.rays
    # shadow rays towards a light behind the surface are counted but not traced (material.zig:55-63)
    assert 0 < full.stats.shadow_rays_culled < full.stats.shadow_rays


@pytest.mark.parametrize("w,h", [(1, 1), (13, 7), (8, 4), (9, 5), (257, 3)])
def test_ragged_frame_sizes(pkg, oracle, w, h):
    """Frames that are not a multiple of the 8x4 work tiles (and the 1x1 frame) against the oracle."""
    compare(pkg, oracle, scenes.SMOKE_SCENE, "", w, h, min_frac=0.99 if w * h > 100 else 1.0, label="ragged")


@pytest.mark.parametrize("depth", [0, 1, 7])
def test_recursion_depths(pkg, oracle, depth):
    """remaining_recursions other than the shipped 5 (camera.zig:118), incl. 0 (no secondary rays at all)."""
    info = pkg.parse_scene(scenes.SMOKE_SCENE, "", 64, 48)
    got = info.camera.render(info.world, max_depth=depth)
    ref = oracle.OracleScene(json_text=scenes.SMOKE_SCENE, width=64, height=48)
    want, counts = ref.render(max_depth=depth)
    assert (np.abs(got.pixels - want).max(axis=2) <= TOL).mean() >= 0.999
    assert got.stats.secondary_rays == counts["secondary"]
    if depth == 0:
        assert got.stats.secondary_rays == 0


def test_empty_world_and_no_lights(pkg, oracle):
    import json
    base = json.loads(scenes.default_world(16, 12))
    empty = dict(base, objects=[])
    info = pkg.parse_scene(json.dumps(empty), "")
    img = info.camera.render(info.world)
    assert not img.pixels.any() and img.stats.primary_rays == 16 * 12 and img.stats.shadow_rays == 0
    dark = dict(base, lights=[])
    info = pkg.parse_scene(json.dumps(dark), "")
    img = info.camera.render(info.world)  # shadeHit sums over zero lights: black, no shadow rays (world.zig:89-96)
    assert not img.pixels.any() and img.stats.shadow_rays == 0


@pytest.mark.parametrize("name", ["smoke", "cover", "fresnel", "dragons"])
def test_megakernel_and_wavefront_are_bit_identical(pkg, name):
    """The two kernels add the same terms weight x lighting() to a pixel as exact integers, so they agree bit for
    bit (and the library may pick whichever is faster for a scene without changing a single pixel)."""
    from tests.golden import make_golden
    if name == "smoke":
        text, assets, (w, h) = scenes.SMOKE_SCENE, "", (160, 96)
    else:
        s = scenes.shipped_scene(name)
        if s is None:
            pytest.skip("reference assets not staged")
        text, assets = s
        w, h = make_golden.SIZES[name]
        w, h = 3 * w, 3 * h
    info = pkg.parse_scene(text, assets, w, h)
    out = {}
    try:
        for k in ("mega", "wave"):
            os.environ["RTC_KERNEL"] = k
            c = info.camera.render(info.world)
            out[k] = (c.pixels.copy(), c.stats.as_dict())
    finally:
        os.environ.pop("RTC_KERNEL", None)
    assert out["mega"][1]["kernel"] == 0 and out["wave"][1]["kernel"] == 1  # megakernel (+ tile order) vs one launch per level
    assert np.array_equal(out["mega"][0], out["wave"][0])
    for key in ("primary", "shadow", "secondary", "shadow_culled"):
        assert out["mega"][1][key] == out["wave"][1][key]


@pytest.mark.parametrize("name", ["smoke", "cover", "dragons", "groups", "align_check"])
def test_lanes_per_pixel_are_bit_identical(pkg, name):
    """render_kernel<MESH, LPP>: 2 or 4 lanes of a warp share a pixel and split its lights; the partial sums are
    integers, so whole frames, partitions, RGBA8 frames and ray counts equal the one-lane-per-pixel kernel's bit
    for bit (the library may use it for small partitions without changing a pixel)."""
    from tests.golden import make_golden
    if name == "smoke":
        text, assets, (w, h) = scenes.SMOKE_SCENE, "", (163, 97)  # ragged: not a multiple of the 8x4 tile
    else:
        s = scenes.shipped_scene(name)
        if s is None:
            pytest.skip("reference assets not staged")
        text, assets = s
        w, h = make_golden.SIZES[name]
        w, h = 2 * w, 2 * h
    info = pkg.parse_scene(text, assets, w, h)
    out = {}
    try:
        os.environ["RTC_KERNEL"] = "mega"
        for lpp in (1, 2, 4, 0):  # 0 = mixed: four lanes on the tiles the previous frame found expensive
            os.environ["RTC_LPP"] = str(lpp)
            # twice each: the second frame of a size is rendered in the order of the first one's tile costs
            # (tile_order.cu), which is what the mixed mode takes its expensive tiles from
            for _ in range(2):
                c = info.camera.render(info.world)
            for _ in range(2):
                p = info.camera.render(info.world, partition=(8, 1, 3))
            for _ in range(2):
                r8, _ = info.camera.render_rgba8(info.world)
            out[lpp] = (c.pixels.copy(), p.pixels.copy(), r8.copy(), c.stats.as_dict())
    finally:
        os.environ.pop("RTC_KERNEL", None)
        os.environ.pop("RTC_LPP", None)
    assert out[1][0].any()
    for lpp in (2, 4, 0):
        for k in range(3):
            assert np.array_equal(out[1][k], out[lpp][k])
        for key in ("primary", "shadow", "secondary", "shadow_culled"):
            assert out[1][3][key] == out[lpp][3][key]


def test_small_partitions_measure_two_lanes_per_pixel(pkg):
    """A partition that leaves the machine underfilled renders frames 5-6 with the mixed kernel (four lanes per pixel
    on the tiles the previous frame found expensive; third candidate of the library's first-frames measurement):
    every frame of the sequence is the same frame."""
    s = scenes.shipped_scene("dragons")
    if s is None:
        pytest.skip("reference assets not staged")
    info = pkg.parse_scene(s[0], s[1], 640, 360)
    frames = [info.camera.render(info.world, partition=(8, 2, 4)).pixels.copy() for _ in range(12)]
    assert frames[0].any()
    for f in frames[1:]:
        assert np.array_equal(frames[0], f)


def test_errors(pkg):
    with pytest.raises(pkg.RtcError):
        pkg.World(b"\0" * 512)
    with pytest.raises(pkg.RtcError):
        pkg.parse_scene("{not json", "")
    info = pkg.parse_scene(scenes.default_world(), "")
    with pytest.raises(pkg.RtcError):
        info.camera.render(info.world, max_depth=99)


@pytest.mark.parametrize("name", ["smoke", "mesh_smooth"] + scenes.SHIPPED)
def test_against_committed_golden(pkg, tmp_path, name):
    """CUDA path vs the committed oracle renders (no oracle code involved on this leg)."""
    from tests.golden import make_golden
    g = np.load(os.path.join(scenes.ROOT, "tests", "golden", "oracle_renders.npz"))
    if name == "smoke":
        text, assets, (w, h) = scenes.SMOKE_SCENE, "", (96, 64)
    elif name == "mesh_smooth":
        scenes.write_mesh_obj(str(tmp_path / "torus.obj"), smooth=True, groups=False)
        text, assets, (w, h) = scenes.mesh_scene("torus.obj"), str(tmp_path), (90, 60)
    else:
        s = scenes.shipped_scene(name)
        if s is None:
            pytest.skip("reference assets not staged (assets/_ref)")
        text, assets, (w, h) = s[0], s[1], make_golden.SIZES[name]
    info = pkg.parse_scene(text, assets, w, h)
    c = info.camera.render(info.world)
    diff = np.abs(c.pixels - g[name]).max(axis=2)
    assert (diff <= TOL).mean() >= 0.999
    st = c.stats
    assert int(st.primary_rays) == int(g[name + "__rays"][0])
    if (diff <= TOL).all():
        assert [int(st.shadow_rays), int(st.secondary_rays)] == [int(v) for v in g[name + "__rays"][1:]]


def test_renderer_twin_of_lib_zig(pkg):
    """include/rtc_renderer.h: the reference's own twelve exports (lib.zig:233-309) driven like www/index.js does —
    init, poll, render, orbit, render again — against the Camera/World mirror on the same scene."""
    import json
    text = scenes.SMOKE_SCENE
    cam_json = json.loads(text)["camera"]
    r = pkg.Renderer(text, ".")
    try:
        assert (r.width, r.height) == (cam_json["width"], cam_json["height"])
        assert not r.pixels.any()  # lib.zig:88-97: zeroed until the first render
        info = pkg.parse_scene(text, "")
        want, st = info.camera.render_rgba8(info.world)
        got = r.render().copy()
        assert np.array_equal(got, want) and (got[..., 3] == 255).all()
        assert r.stats().rays == st.rays
        # lib.zig:166-190 with numpy: orbit about `up` through `to`, then dolly towards `to`
        frm, to, up = (np.array(cam_json[k], dtype=np.float64) for k in ("from", "to", "up"))
        angle, dist = 0.35, 0.2
        C = np.array([[0, -up[2], up[1]], [up[2], 0, -up[0]], [-up[1], up[0], 0]])
        R = np.eye(3) + C * np.sin(angle) + (C @ C) * (1 - np.cos(angle))
        frm = R @ (frm - to) + to
        frm = frm + (to - frm) * dist
        r.rotate_camera(angle)
        r.move_camera(dist)
        moved = r.render().copy()
        assert not np.array_equal(moved, got)
        info.camera.look_at(frm, to, up)
        want2, _ = info.camera.render_rgba8(info.world)
        diff = np.abs(moved.astype(int) - want2.astype(int))
        assert (diff <= 1).mean() > 0.999  # `from` differs in the last bits between numpy and the C++ mirror
    finally:
        r.close()


def parity_record(name, w, h, canvas, want, counts, extra=None):
    """One line of the parity report (profiles/parity_r2.jsonl): the north star's three criteria on the WHOLE frame."""
    diff = np.abs(canvas.pixels - want).max(axis=2)
    ok = diff <= TOL
    q = np.abs(canvas.to_image().astype(int) - quantise_ref(want).astype(int)).max(axis=2)
    st = canvas.stats
    ys, xs = np.nonzero(~ok)
    rec = {"scene": name, "width": w, "height": h, "pixels": int(ok.size), "frac_within_1e-4": float(ok.mean()),
           "divergent_pixels": int((~ok).sum()), "divergent_xy": [[int(x), int(y)] for x, y in zip(xs[:32], ys[:32])],
           "max_abs_diff": float(diff.max()), "max_abs_diff_on_ok": float(diff[ok].max()) if ok.any() else 0.0,
           "max_8bit_diff": int(q.max()), "max_8bit_diff_on_ok": int(q[ok].max()) if ok.any() else 0,
           "rays_gpu": [int(st.primary_rays), int(st.shadow_rays), int(st.secondary_rays)],
           "rays_oracle": [int(counts["primary"]), int(counts["shadow"]), int(counts["secondary"])],
           "kernel_ms": float(st.kernel_ms)}
    rec.update(extra or {})
    return rec, ok, q


def quantise_ref(rgb):
    from oracle import oracle_py
    return oracle_py.quantise(rgb)


def write_parity(rec):
    out = os.path.join(scenes.ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, "parity_r2.jsonl"), "a") as f:
        f.write(json.dumps(rec) + "\n")


def check_parity(rec, ok, q):
    assert rec["frac_within_1e-4"] >= 0.999
    assert q[ok].max() <= 1
    assert rec["rays_gpu"][0] == rec["rays_oracle"][0] == rec["pixels"]
    if rec["divergent_pixels"] == 0:
        assert rec["rays_gpu"] == rec["rays_oracle"]  # P, S, Q exactly


@pytest.mark.parametrize("name,w,h", [("fresnel", 600, 600), ("cover", 1920, 1080), ("teapot", 1920, 1080),
                                      ("dragons", 1920, 1080), ("csg", 3840, 2160), ("skybox", 3840, 2160)])
def test_baseline_config_full_size(pkg, oracle, name, w, h):
    """BASELINE.json's configs at their FULL sizes, WHOLE frame against the oracle's whole frame (the oracle needs a few
    seconds for dragons on the box's cores, less for the others): >= 99.9 % of the pixels within 1e-4, 8-bit within
    +-1 LSB, P/S/Q equal when no pixel diverges; plus the size-independent properties -- the union of three interleaved
    partitions is the frame, the RGBA8 twin is clamp() of the f64 frame, the ray counts of the partitions add up."""
    s = scenes.shipped_scene(name)
    if s is None:
        pytest.skip("reference assets not staged (assets/_ref)")
    text, assets = s
    info = pkg.parse_scene(text, assets, w, h)
    canvas = info.camera.render(info.world)
    assert np.isfinite(canvas.pixels).all() and canvas.stats.primary_rays == w * h
    ref = oracle.OracleScene(json_text=text, asset_dir=assets, width=w, height=h)
    want, counts = ref.render()
    rec, ok, q = parity_record(name, w, h, canvas, want, counts, {"config": "BASELINE.json"})
    print(json.dumps(rec))
    write_parity(rec)
    check_parity(rec, ok, q)
    acc = np.zeros_like(canvas.pixels)
    rays = 0
    for part in range(3):
        c = info.camera.render(info.world, partition=(8, part, 3), out=acc)
        rays += c.stats.rays
    assert np.array_equal(acc, canvas.pixels) and rays == canvas.stats.rays
    rgba, _ = info.camera.render_rgba8(info.world)
    assert np.array_equal(rgba[..., :3], oracle.quantise(canvas.pixels)) and (rgba[..., 3] == 255).all()
    info.world.close()
    ref.close()


@pytest.mark.parametrize("name", scenes.SHIPPED)
def test_shipped_scene_native_resolution(pkg, oracle, name):
    """Every shipped scene (main.zig:32-47) at the resolution its JSON asks for, whole frame: the north star's
    correctness clause, one report line per scene."""
    s = scenes.shipped_scene(name)
    if s is None:
        pytest.skip("reference assets not staged (assets/_ref)")
    text, assets = s
    info = pkg.parse_scene(text, assets)
    w, h = info.camera.hsize, info.camera.vsize
    canvas = info.camera.render(info.world)
    ref = oracle.OracleScene(json_text=text, asset_dir=assets)
    want, counts = ref.render()
    rec, ok, q = parity_record(name, w, h, canvas, want, counts, {"config": "native"})
    print(json.dumps(rec))
    write_parity(rec)
    check_parity(rec, ok, q)
    info.world.close()
    ref.close()


def test_async_frames_and_stats_fetch(pkg):
    """rtc_render_device_async only puts frames on the stream; rtc_stats_fetch hands out their statistics in order.
    Same frame, same counts as the synchronous call; frames of one scene put on DIFFERENT streams stay ordered
    (the tile-order kernels of frame k must not overlap the memsets of frame k + 1)."""
    import torch
    s = scenes.shipped_scene("teapot")
    text, assets, (w, h) = (s[0], s[1], (400, 240)) if s else (scenes.SMOKE_SCENE, "", (160, 96))
    info = pkg.parse_scene(text, assets, w, h)
    cam, wld = info.camera, info.world
    ref = torch.empty((h, w, 3), dtype=torch.float64, device="cuda")
    for _ in range(7):  # past the library's first-frames measurement
        want = cam.render_device_frame(wld, ref.data_ptr())
    frames = [torch.full((h, w, 3), float("nan"), dtype=torch.float64, device="cuda") for _ in range(6)]
    side = torch.cuda.Stream()
    for i, f in enumerate(frames):
        stream = side.cuda_stream if i % 2 else torch.cuda.current_stream().cuda_stream
        cam.render_device_async(wld, f.data_ptr(), stream=stream)
    stats = wld.fetch_stats()
    torch.cuda.synchronize()
    assert len(stats) == len(frames)
    for f, st in zip(frames, stats):
        assert torch.equal(f, ref)
        assert (st.primary_rays, st.shadow_rays, st.secondary_rays) == (want.primary_rays, want.shadow_rays, want.secondary_rays)
        assert st.kernel_ms > 0
    assert wld.fetch_stats() == []
    # a synchronous call after un-fetched asynchronous ones returns ITS OWN statistics
    cam.render_device_async(wld, frames[0].data_ptr())
    cam.look_at((0, 3, -8), (0, 1, 0), (0, 1, 0))
    moved = cam.render_device_frame(wld, frames[1].data_ptr())
    assert moved.primary_rays == w * h and not torch.equal(frames[1], ref)
    wld.close()


@pytest.mark.parametrize("name", ["dragons", "teapot", "cover", "fresnel", "xyz", "cubes", "reflection_and_refraction", "csg",
                                  "skybox", "nefertiti", "mesh"])
def test_lean_and_general_kernels_are_bit_identical(pkg, tmp_path, name):
    """render_kernel<MESH, LPP, FEAT>: a scene runs the instantiation that carries only the machinery it needs (no CSG /
    guard code without CSG objects; the inline simple_pattern_color instead of the general pattern evaluator where
    every pattern is a solid colour or one simple pattern over two); RTC_LEAN=0 keeps the general one (FEAT = 3).
    Same frame, same ray counts, bit for bit."""
    if name == "mesh":
        scenes.write_mesh_obj(str(tmp_path / "torus.obj"), smooth=True, groups=False)
        text, assets, (w, h) = scenes.mesh_scene("torus.obj"), str(tmp_path), (200, 120)
    else:
        s = scenes.shipped_scene(name)
        if s is None:
            pytest.skip("reference assets not staged")
        text, assets, (w, h) = s[0], s[1], (320, 180)
    out = {}
    try:
        for key, lean in (("lean", "1"), ("general", "0")):
            os.environ["RTC_LEAN"] = lean  # read when the scene is created
            info = pkg.parse_scene(text, assets, w, h)
            c = info.camera.render(info.world)
            out[key] = (c.pixels.copy(), (c.stats.primary_rays, c.stats.shadow_rays, c.stats.secondary_rays))
            info.world.close()
    finally:
        os.environ.pop("RTC_LEAN", None)
    assert out["lean"][0].any()
    assert np.array_equal(out["lean"][0], out["general"][0]) and out["lean"][1] == out["general"][1]


def test_two_scenes_keep_their_shared_memory(pkg, tmp_path):
    """the dynamic shared memory limit belongs to the kernel, not to a scene -- a deep scene created first,
    a shallow one second, then the deep one rendered again."""
    scenes.write_mesh_obj(str(tmp_path / "torus.obj"), smooth=True, groups=False)
    deep = pkg.parse_scene(scenes.mesh_scene("torus.obj"), str(tmp_path), 96, 64)
    a = deep.camera.render(deep.world).pixels.copy()
    shallow = pkg.parse_scene(scenes.default_world(), "")
    shallow.camera.render(shallow.world)
    b = deep.camera.render(deep.world).pixels
    assert np.array_equal(a, b)


def test_rgba8_into_pinned_memory_is_zero_copy_and_equal(pkg):
    """rtc_render_rgba8 into rtc_host_alloc memory (what the lib.zig twin's pixel buffer is): the kernel writes it
    directly; same bytes as the staged path, whole frame and partitions."""
    info = pkg.parse_scene(scenes.SMOKE_SCENE, "", 163, 97)
    staged, _ = info.camera.render_rgba8(info.world)
    pinned = pkg.pinned_empty((97, 163, 4), dtype=np.uint8)
    pinned[:] = 0
    st = pkg.rtc_stats()
    import ctypes
    for part in (None, (8, 0, 2), (8, 1, 2)):
        keep, pref = info.camera._part(part)
        if part == (8, 0, 2):
            pinned[:] = 0
        pkg._check(pkg.lib().rtc_render_rgba8(info.world._h, ctypes.byref(info.camera.c), 5, pref, pinned.ctypes.data, ctypes.byref(st)))
        if part is None or part == (8, 1, 2):
            assert np.array_equal(pinned, staged)
    part_only, _ = info.camera.render_rgba8(info.world, partition=(8, 1, 3))  # staged + strided 2-D copy
    rows = [y for y in range(97) if (y // 8) % 3 == 1]
    assert np.array_equal(part_only[rows], staged[rows])
    other = [y for y in range(97) if (y // 8) % 3 != 1]
    assert not part_only[other].any()
    pkg.free_pinned(pinned)


@pytest.mark.parametrize("w,h", [(163, 97), (7, 5), (1, 1), (640, 363)])
def test_pinned_canvas_through_the_copy_engine(pkg, w, h):
    """rtc_render / rtc_render_rgba8 into rtc_host_alloc memory: the kernel announces finished rows of tiles and the calling
    thread hands them to the copy engine while the rest renders (rtc_api.cu render_to_pinned).  Same frame as the staged
    path bit for bit -- ragged sizes, the first frames (no order yet, kernel measurement, wavefront frames copied
    afterwards), a camera that moves every frame (new band order every frame), f64 and RGBA8 alternating."""
    import ctypes
    info = pkg.parse_scene(scenes.SMOKE_SCENE, "", w, h)
    cam, wld = info.camera, info.world
    pinned = pkg.pinned_empty((h, w, 3))
    pin8 = pkg.pinned_empty((h, w, 4), dtype=np.uint8)
    st = pkg.rtc_stats()
    for k in range(14):
        if k >= 8:
            cam.look_at((0.3 * (k - 8), 2.5, -7), (0, 0.8, 0), (0, 1, 0))
        want = cam.render(wld).pixels          # pageable canvas: staged, one copy afterwards
        want8, _ = cam.render_rgba8(wld)
        pinned[:] = np.nan
        got = cam.render(wld, out=pinned)
        assert np.array_equal(pinned, want), (w, h, k)
        assert got.stats.primary_rays == w * h
        pin8[:] = 0
        pkg._check(pkg.lib().rtc_render_rgba8(wld._h, ctypes.byref(cam.c), 5, None, pin8.ctypes.data, ctypes.byref(st)))
        assert np.array_equal(pin8, want8), (w, h, k)
    pkg.free_pinned(pinned)
    pkg.free_pinned(pin8)
    wld.close()


def test_replicated_scene_uses_all_devices(pkg):
    """rtc_set_devices(n): ONE process, the scene built once and replicated, a whole-frame call split over the devices
    (interleaved 8-row tiles, peer stores into device 0's frame / into the caller's pinned canvas): bit-identical."""
    import torch
    if pkg.device_count() < 2:
        pytest.skip("one GPU visible")
    n = min(pkg.device_count(), 4)
    s = scenes.shipped_scene("dragons")
    text, assets, (w, h) = (s[0], s[1], (480, 270)) if s else (scenes.SMOKE_SCENE, "", (160, 96))
    blob, cam = pkg.load_blob(text, assets, w, h)
    one = pkg.World(blob, 0)
    camera = pkg.Camera(cam)
    want = camera.render(one)
    many = pkg.World(blob, 0, devices=n)
    got = camera.render(many)                              # pageable canvas: assembled on device 0, one copy
    assert np.array_equal(got.pixels, want.pixels) and got.stats.rays == want.stats.rays
    pinned = pkg.pinned_empty((h, w, 3))
    got2 = camera.render(many, out=pinned)                  # pinned canvas: every device stores its rows itself
    assert np.array_equal(got2.pixels, want.pixels)
    frame = torch.zeros((h, w, 3), dtype=torch.float64, device="cuda:0")
    st = camera.render_device_frame(many, frame.data_ptr())
    torch.cuda.synchronize()
    assert np.array_equal(frame.cpu().numpy(), want.pixels) and st.rays == want.stats.rays
    rgba, _ = camera.render_rgba8(many)
    assert np.array_equal(rgba[..., :3], want.to_image())
    pkg.free_pinned(pinned)
    many.close()
    one.close()


@pytest.mark.parametrize("n", [2, 3])
def test_frame_shared_tile_by_tile_is_the_same_frame(pkg, n):
    """rtc_scene_share_tiles on ONE GPU: n scenes (the "ranks") take the tiles of one cost order in turn and add every
    tile's cost to all n cost buffers.  Rendered round by round into one frame, every round gives the whole frame, bit
    for bit, and the ray counts add up -- first round (no order yet), later rounds (cost order, mixed kernel included)."""
    import torch
    s = scenes.shipped_scene("dragons")
    text, assets, (w, h) = (s[0], s[1], (500, 203)) if s else (scenes.SMOKE_SCENE, "", (163, 97))
    blob, cam = pkg.load_blob(text, assets, w, h)
    camera = pkg.Camera(cam)
    ref_world = pkg.World(blob, 0)
    want = camera.render(ref_world)
    words = int(pkg.lib().rtc_tile_share_words(w, h))
    bufs = [torch.zeros(2 * words, dtype=torch.int32, device="cuda") for _ in range(n)]
    ptrs = [[bufs[r].data_ptr() + b * words * 4 for r in range(n)] for b in range(2)]
    worlds = [pkg.World(blob, 0) for _ in range(n)]
    for r, wld in enumerate(worlds):
        wld.share_tiles(r, n, words, ptrs)
    frame = torch.empty((h, w, 3), dtype=torch.float64, device="cuda")
    for rnd in range(12):  # past the first-frames measurement: rounds 4-9 run the mixed kernel on small shares
        frame.fill_(float("nan"))
        rays = 0
        for wld in worlds:
            rays += camera.render_device_frame(wld, frame.data_ptr()).rays
        torch.cuda.synchronize()
        assert np.array_equal(frame.cpu().numpy(), want.pixels), "round %d" % rnd
        assert rays == want.stats.rays
    # every rank's copy of the costs is the same
    torch.cuda.synchronize()
    for r in range(1, n):
        assert torch.equal(bufs[0], bufs[r])
    worlds[0].share_tiles(0, 1, 0, None)  # back to a renderer by itself
    frame.fill_(0)
    camera.render_device_frame(worlds[0], frame.data_ptr())
    torch.cuda.synchronize()
    assert np.array_equal(frame.cpu().numpy(), want.pixels)


@pytest.mark.parametrize("name", ["teapot", "dragons", "nefertiti", "mesh"])
def test_gpu_built_and_host_built_bvh_give_the_same_frame(pkg, tmp_path, name):
    """the BVH of a mesh built on the device (LBVH, bvh_gpu.cu; RTC_BVH=gpu) and on the host (binned
    SAH, bvh.cpp; RTC_BVH=host).  Both trees only cull with conservative boxes, hits and ties are decided by the f64
    test and the leaf order: the same frame and the same ray counts, bit for bit."""
    if name == "mesh":
        scenes.write_mesh_obj(str(tmp_path / "torus.obj"), smooth=True, groups=False)
        text, assets, (w, h) = scenes.mesh_scene("torus.obj"), str(tmp_path), (200, 120)
    else:
        s = scenes.shipped_scene(name)
        if s is None:
            pytest.skip("reference assets not staged")
        text, assets, (w, h) = s[0], s[1], (320, 200)
    out = {}
    try:
        for mode in ("host", "gpu"):
            os.environ["RTC_BVH"] = mode  # read when the scene is created
            info = pkg.parse_scene(text, assets, w, h)
            c = info.camera.render(info.world)
            out[mode] = (c.pixels.copy(), (c.stats.primary_rays, c.stats.shadow_rays, c.stats.secondary_rays))
            info.world.close()
    finally:
        os.environ.pop("RTC_BVH", None)
    assert out["host"][0].any()
    assert np.array_equal(out["host"][0], out["gpu"][0]) and out["host"][1] == out["gpu"][1]
