"""CPU-side checks: the C-ABI library loads and exports every declared symbol, the host loader +
flattener produce a well-formed blob, and compute entry points fail loudly without a GPU."""
import ctypes
import json
import re
import struct

import numpy as np
import pytest

from tests import scenes


def test_library_exports_every_declared_symbol(pkg):
    import os
    hdr = open(os.path.join(scenes.ROOT, "include", "rtc_b200.h")).read()
    declared = set(re.findall(r"\b(rtc_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(pkg.EXPORTS)
    L = pkg.lib()
    for name in declared:
        assert hasattr(L, name), name
    # include/rtc_renderer.h: the reference's own twelve exports (lib.zig:233-309), same names
    hdr = open(os.path.join(scenes.ROOT, "include", "rtc_renderer.h")).read()
    hdr = hdr[hdr.index('extern "C"'):]
    declared = set(re.findall(r"\b([a-zA-Z]+)\s*\((?:void|const|double|size_t)", hdr))
    assert declared == set(pkg.RENDERER_EXPORTS), declared ^ set(pkg.RENDERER_EXPORTS)
    for name in ("startInitRenderer", "tryFinishInitRenderer", "initRendererIsOk", "initRendererGetPixels",
                 "initRendererGetWidth", "initRendererGetHeight", "initRendererGetErr", "deinitRenderer",
                 "startRender", "tryFinishRender", "rotateCamera", "moveCamera"):
        assert name in declared and hasattr(L, name), name


def test_renderer_init_reports_errors(pkg):
    """lib.zig:216-231: a scene that does not parse ends in initRendererIsOk() == false + a message, not a crash."""
    with pytest.raises(pkg.RtcError) as e:
        pkg.Renderer("{not json")
    assert str(e.value)


def parse_header(blob):
    magic, version, total = struct.unpack_from("<IIQ", blob, 0)
    counts = struct.unpack_from("<12I", blob, 16)
    return magic, version, total, counts


def test_host_loader_blob(pkg):
    blob, cam = pkg.load_blob(scenes.SMOKE_SCENE, "", 96, 64)
    magic, version, total, counts = parse_header(blob)
    assert magic == 0x42435452 and version == 2 and total == len(blob) and len(blob) % 16 == 0
    n_lights, n_mat, n_pat, n_uv, n_img, n_xf, n_leaves, n_tris, n_guards, n_nodes, n_children, n_csg = counts
    assert n_lights == 2
    assert n_leaves == 11 and n_tris == 1      # plane, 2 spheres, cube, cylinder, cone, csg(2), group(3)
    assert n_csg == 1 and n_nodes == 3         # csg node + its two leaf nodes
    assert n_guards == 0
    assert (cam.hsize, cam.vsize) == (96, 64)
    # camera.zig:33-52 : pixel size for a 96x64 frame, fov 0.9
    import math
    half = math.tan(0.45)
    assert abs(cam.half_width - half) < 1e-15 and abs(cam.pixel_size - 2 * half / 96) < 1e-15


def test_resolution_override_matches_json_edit(pkg):
    a, cam_a = pkg.load_blob(scenes.SMOKE_SCENE, "", 40, 30)
    b, cam_b = pkg.load_blob(scenes.with_size(scenes.SMOKE_SCENE, 40, 30), "")
    assert a == b and bytes(cam_a) == bytes(cam_b)


def test_unbounded_child_in_group_needs_no_guard(pkg):
    # A plane inside a group: its parentSpaceBounds are all NaN (0 * inf), merging them makes the
    # group box (-inf,+inf), and the group's own setTransform turns that into the untouched box
    # (bounding_box.zig:24-32,49-70; shape.zig:294-296).  Both never cull, so no guard is emitted.
    scene = json.loads(scenes.default_world())
    scene["objects"] = [{"type": {"group": [{"type": {"sphere": {}}}, {"type": {"plane": {}}}]}}]
    blob, _ = pkg.load_blob(json.dumps(scene), "")
    counts = parse_header(blob)[3]
    assert counts[6] == 2 and counts[8] == 0  # leaves, guards


def test_mesh_blob(pkg, tmp_path):
    scenes.write_mesh_obj(str(tmp_path / "torus.obj"), n_u=12, n_v=8, smooth=True)
    blob, _ = pkg.load_blob(scenes.mesh_scene("torus.obj"), str(tmp_path))
    counts = parse_header(blob)[3]
    assert counts[7] == 2 * 12 * 8 * 2  # two instances, quads fan-triangulated
    assert counts[5] >= 4               # plane, 2 mesh transforms, cube


def test_parse_errors(pkg):
    with pytest.raises(pkg.RtcError) as e:
        pkg.load_blob("{", "")
    assert e.value.code == pkg.RTC_ERR_PARSE
    with pytest.raises(pkg.RtcError):
        pkg.load_blob(json.dumps({"camera": {}, "lights": [], "objects": []}), "")


def test_partition_rows(pkg):
    L = pkg.lib()
    for vsize in (1, 7, 64, 1080):
        for tile in (1, 4, 8, 16):
            for n in (1, 2, 3, 8):
                tot = 0
                for p in range(n):
                    part = pkg.rtc_partition(tile, p, n)
                    tot += L.rtc_partition_rows(vsize, ctypes.byref(part))
                assert tot == vsize


def test_compute_fails_loudly_without_gpu(pkg):
    if pkg.device_count() > 0:
        pytest.skip("a GPU is present")
    blob, cam = pkg.load_blob(scenes.default_world(), "")
    with pytest.raises(pkg.RtcError) as e:
        pkg.World(blob)
    assert e.value.code == pkg.RTC_ERR_CUDA


def test_oracle_default_world_kat(oracle):
    # camera.zig:171-187
    s = oracle.OracleScene(json_text=scenes.default_world())
    img, counts = s.render(threads=2)
    assert np.allclose(img[5, 5], [0.38066, 0.47583, 0.2855], atol=1e-5)
    assert counts["primary"] == 121


def test_canvas_ppm_golden(pkg):
    """canvas.zig:259-301: the reference's own two PPM vectors (clamping, and lines broken before column 70)."""
    px = np.zeros((3, 5, 3))
    px[0, 0, 0] = 1.5
    px[1, 2, 1] = 0.5
    px[2, 4] = (-0.5, 0.0, 1.0)
    assert pkg.Canvas(px).ppm() == ("P3\n5 3\n255\n"
                                    "255 0 0 0 0 0 0 0 0 0 0 0 0 0 0\n"
                                    "0 0 0 0 0 0 0 128 0 0 0 0 0 0 0\n"
                                    "0 0 0 0 0 0 0 0 0 0 0 0 0 0 255\n")
    px = np.empty((2, 10, 3))
    px[:] = (1, 0.8, 0.6)
    row = ("255 204 153 255 204 153 255 204 153 255 204 153 255 204 153 255 204\n"
           "153 255 204 153 255 204 153 255 204 153 255 204 153\n")
    assert pkg.Canvas(px).ppm() == "P3\n10 2\n255\n" + row + row


def test_png_round_trip(pkg, tmp_path):
    """rtc_write_png (main.zig:95-99's rgb24 PNG) read back with nothing but zlib."""
    import zlib
    rng = np.random.default_rng(5)
    px = rng.random((7, 13, 3)) * 1.2 - 0.1
    c = pkg.Canvas(px)
    path = tmp_path / "out.png"
    c.save_png(path)
    data = path.read_bytes()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, ihdr = 8, b"", None
    while pos < len(data):
        n, kind = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        assert struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])[0] == zlib.crc32(kind + body)
        if kind == b"IHDR":
            ihdr = struct.unpack(">IIBBBBB", body)
        elif kind == b"IDAT":
            idat += body
        pos += 12 + n
    assert ihdr == (13, 7, 8, 2, 0, 0, 0)
    raw = np.frombuffer(zlib.decompress(idat), dtype=np.uint8).reshape(7, 1 + 13 * 3)
    assert not raw[:, 0].any()
    assert np.array_equal(raw[:, 1:].reshape(7, 13, 3), c.to_image())
