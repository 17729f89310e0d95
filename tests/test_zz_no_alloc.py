"""rtc_render_device_static: the render call that allocates nothing (scratch in a fixed __device__ array of the library).

Collected LAST on purpose.  Through the C ABI, against the existing entry point
rtc_render_device on the same scene and camera; the tolerance is the one the tests use for two runs of the same kernel
arithmetic: none, frames are bit-identical (every pixel is summed by itself in exact integers, whatever the tile order)."""
import pytest

from tests import scenes

pytestmark = pytest.mark.gpu


def camera_copy(pkg, cam):
    return pkg.Camera(type(cam.c).from_buffer_copy(cam.c))


def frames_of(pkg, tmp_path):
    """(label, scene text, asset dir, width, height): an analytic scene, a triangle mesh, another frame size."""
    scenes.write_mesh_obj(str(tmp_path / "torus.obj"), smooth=True, groups=False)
    out = [("smoke", scenes.SMOKE_SCENE, "", 192, 128), ("mesh", scenes.mesh_scene("torus.obj"), str(tmp_path), 200, 120)]
    s = scenes.shipped_scene("teapot")
    if s is not None:
        out.append(("teapot", s[0], s[1], 400, 240))
    return out


def test_static_scratch_render_equals_the_allocating_entry_point(pkg, tmp_path):
    import torch
    L = pkg.lib()
    assert L.rtc_static_max_tiles() == 1 << 18
    worlds = []
    for label, text, assets, w, h in frames_of(pkg, tmp_path):
        info = pkg.parse_scene(text, assets, w, h)
        cam, wld = info.camera, info.world
        want = torch.empty((h, w, 3), dtype=torch.float64, device="cuda")
        st_want = cam.render_device(wld, want.data_ptr())  # the existing entry point
        got = torch.empty_like(want)
        worlds.append((label, cam, wld, want, st_want, got))
    # two calls in a row per scene (the second runs in the tile order made from the first one's costs), then the scenes
    # in turn: every call finds the array as ANOTHER scene and frame size left it
    for rnd in range(3):
        for label, cam, wld, want, st_want, got in worlds:
            for call in range(2 if rnd == 0 else 1):
                got.fill_(float("nan"))
                st = cam.render_device_static(wld, got.data_ptr())
                assert torch.equal(got, want), (label, rnd, call)
                assert (st.primary_rays, st.shadow_rays, st.secondary_rays, st.shadow_rays_culled) == \
                    (st_want.primary_rays, st_want.shadow_rays, st_want.secondary_rays, st_want.shadow_rays_culled), (label, rnd, call)
                assert st.pixels == cam.hsize * cam.vsize and st.kernel_ms > 0
                assert st.launches == (3 if (rnd == 0 and call == 1) else 1), (label, rnd, call, st.launches)
    # enqueue only (stats = NULL), on a side stream, frames back to back; a moved camera; no device memory is taken
    label, cam, wld, want, st_want, got = worlds[-1]
    side = torch.cuda.Stream()
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    for _ in range(3):
        got.fill_(float("nan"))
        torch.cuda.synchronize()
        assert cam.render_device_static(wld, got.data_ptr(), stream=side.cuda_stream, wait=False) is None
        side.synchronize()
        assert torch.equal(got, want)
    moved = camera_copy(pkg, cam)
    moved.look_at((0.5, 3, -8), (0, 1, 0), (0, 1, 0))
    for _ in range(3):
        st = moved.render_device_static(wld, got.data_ptr())
        assert st.primary_rays == cam.hsize * cam.vsize and not torch.equal(got, want)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free_before
    moved.render_device(wld, want.data_ptr())
    assert torch.equal(got, want)
    for _, _, wld, _, _, _ in worlds:
        wld.close()


def test_static_scratch_refuses_frames_beyond_its_size(pkg):
    import torch
    info = pkg.parse_scene(scenes.SMOKE_SCENE, "", 64, 32)
    cam, wld = info.camera, info.world
    small = torch.zeros((32, 64, 3), dtype=torch.float64, device="cuda")
    big = camera_copy(pkg, cam)
    big.c.hsize, big.c.vsize = 8192, 4096  # 1,048,576 tiles of 8x4 > 262,144: refused before anything is touched
    with pytest.raises(pkg.RtcError) as e:
        big.render_device_static(wld, small.data_ptr())
    assert e.value.code == pkg.RTC_ERR_INVALID and "static scratch holds 262144" in str(e.value)
    edge = camera_copy(pkg, cam)
    edge.c.hsize, edge.c.vsize = 8 * 512 + 1, 4 * 511 + 1  # 513 x 512 tiles = 262,656: just beyond
    with pytest.raises(pkg.RtcError):
        edge.render_device_static(wld, small.data_ptr())
    with pytest.raises(pkg.RtcError):
        cam.render_device_static(wld, 0)
    st = cam.render_device_static(wld, small.data_ptr())  # the refusals left nothing behind
    assert st.primary_rays == 64 * 32
    wld.close()
