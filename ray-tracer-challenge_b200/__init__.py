"""B200-native renderer for the per-pixel hot path of SinclaM/ray-tracer-challenge.

Host-side mirror of the reference's operator interface for this path:

    info   = parse_scene(json_text, asset_dir)         # src/parsing/scene.zig:612  parseScene
    canvas = info.camera.render(info.world)            # src/raytracer/camera.zig:80 Camera.render
    canvas.pixels          # [H, W, 3] float64 linear RGB, row-major, y down (canvas.zig:18-21)
    canvas.to_image()      # [H, W, 3] uint8, clamp() quantised       (canvas.zig:150-176)

Everything is a thin ctypes layer over the C ABI in include/rtc_b200.h (librtc_b200.so, built
in-tree by `make` / __graft_entry__.build()).  There is no CPU fallback: importing works anywhere,
but creating a World needs the library and a CUDA device, and fails loudly otherwise.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# RTC_B200_LIB: an alternative build of the SAME library (kernel tuning experiments, tools/)
LIB_PATH = os.environ.get("RTC_B200_LIB") or os.path.join(_HERE, "librtc_b200.so")
_lib = None

RTC_OK, RTC_ERR_INVALID, RTC_ERR_CUDA, RTC_ERR_PARSE, RTC_ERR_CAPACITY = 0, -1, -2, -3, -4


class RtcError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("rtc_b200 error %d: %s" % (code, message))
        self.code = code


class rtc_camera(ctypes.Structure):
    _fields_ = [("hsize", ctypes.c_uint64), ("vsize", ctypes.c_uint64), ("half_width", ctypes.c_double),
                ("half_height", ctypes.c_double), ("pixel_size", ctypes.c_double), ("inverse", ctypes.c_double * 16)]


class rtc_partition(ctypes.Structure):
    _fields_ = [("tile_rows", ctypes.c_uint32), ("part", ctypes.c_uint32), ("n_parts", ctypes.c_uint32)]


class rtc_scene_info(ctypes.Structure):
    _fields_ = [("n_prims", ctypes.c_uint32), ("n_meshes", ctypes.c_uint32), ("n_geometries", ctypes.c_uint32),
                ("n_csg", ctypes.c_uint32), ("n_bvh_nodes", ctypes.c_uint64), ("n_mesh_triangles", ctypes.c_uint64),
                ("device_bytes", ctypes.c_uint64), ("bvh_max_depth", ctypes.c_uint32), ("n_leaves", ctypes.c_uint32),
                ("build_seconds", ctypes.c_double)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class rtc_tile_share(ctypes.Structure):
    _fields_ = [("rank", ctypes.c_uint32), ("n_ranks", ctypes.c_uint32), ("words", ctypes.c_uint64),
                ("cost", (ctypes.c_void_p * 8) * 2)]


class rtc_stats(ctypes.Structure):
    _fields_ = [("primary_rays", ctypes.c_uint64), ("shadow_rays", ctypes.c_uint64),
                ("secondary_rays", ctypes.c_uint64), ("pixels", ctypes.c_uint64), ("kernel_ms", ctypes.c_float),
                ("kernel", ctypes.c_uint32), ("launches", ctypes.c_uint64), ("shadow_rays_culled", ctypes.c_uint64)]

    def as_dict(self):
        return {"primary": self.primary_rays, "shadow": self.shadow_rays, "secondary": self.secondary_rays,
                "pixels": self.pixels, "kernel_ms": self.kernel_ms, "launches": self.launches, "kernel": self.kernel,
                "shadow_culled": self.shadow_rays_culled}

    @property
    def rays(self):
        return self.primary_rays + self.shadow_rays + self.secondary_rays


# every symbol include/rtc_b200.h declares
EXPORTS = ["rtc_last_error", "rtc_device_count", "rtc_set_device", "rtc_set_devices", "rtc_render_device_async", "rtc_stats_fetch", "rtc_tile_share_words", "rtc_scene_share_tiles", "rtc_blob_inspect", "rtc_scene_create", "rtc_scene_destroy",
           "rtc_render", "rtc_render_rgba8", "rtc_render_device", "rtc_render_device_frame", "rtc_partition_rows", "rtc_load_scene_json",
           "rtc_blob_free", "rtc_camera_look_at", "rtc_host_alloc", "rtc_host_free", "rtc_canvas_ppm", "rtc_write_png",
           "rtc_static_max_tiles", "rtc_render_device_static"]
# include/rtc_renderer.h: the twelve exports of the reference's own library (src/lib.zig:233-309) + two of ours
RENDERER_EXPORTS = ["startInitRenderer", "tryFinishInitRenderer", "initRendererIsOk", "initRendererGetPixels",
                    "initRendererGetWidth", "initRendererGetHeight", "initRendererGetErr", "deinitRenderer",
                    "startRender", "tryFinishRender", "rotateCamera", "moveCamera", "rtcRendererSetAssetDir",
                    "rtcRendererLastStats"]


def lib():
    """Loads librtc_b200.so (no compute happens at load time)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RtcError(RTC_ERR_CUDA, "%s is missing: run `make` (or __graft_entry__.build()); there is no fallback path"
                       % LIB_PATH)
    L = ctypes.CDLL(LIB_PATH)
    vp, cp = ctypes.c_void_p, ctypes.c_char_p
    L.rtc_last_error.restype = cp
    L.rtc_device_count.restype = ctypes.c_int
    L.rtc_set_device.argtypes = [ctypes.c_int]
    if not hasattr(L, "rtc_set_devices"):  # an older build of the library (tools/kernel_variants.py compares against one)
        return _finish_old(L)
    L.rtc_set_devices.argtypes = [ctypes.c_int]
    L.rtc_render_device_async.argtypes = [vp, ctypes.POINTER(rtc_camera), ctypes.c_int, ctypes.POINTER(rtc_partition), vp, vp]
    L.rtc_stats_fetch.argtypes = [vp, ctypes.POINTER(rtc_stats), ctypes.c_uint32, ctypes.POINTER(ctypes.c_uint32)]
    L.rtc_tile_share_words.argtypes = [ctypes.c_uint64, ctypes.c_uint64]
    L.rtc_tile_share_words.restype = ctypes.c_uint64
    L.rtc_scene_share_tiles.argtypes = [vp, ctypes.POINTER(rtc_tile_share)]
    L.rtc_blob_inspect.argtypes = [vp, ctypes.c_size_t, ctypes.POINTER(rtc_scene_info)]
    L.rtc_scene_create.argtypes = [vp, ctypes.c_size_t, ctypes.POINTER(vp)]
    L.rtc_scene_destroy.argtypes = [vp]
    L.rtc_scene_destroy.restype = None
    common = [vp, ctypes.POINTER(rtc_camera), ctypes.c_int, ctypes.POINTER(rtc_partition)]
    L.rtc_render.argtypes = common + [vp, ctypes.POINTER(rtc_stats)]
    L.rtc_render_rgba8.argtypes = common + [vp, ctypes.POINTER(rtc_stats)]
    L.rtc_render_device.argtypes = common + [vp, vp, ctypes.POINTER(rtc_stats)]
    L.rtc_render_device_frame.argtypes = common + [vp, vp, ctypes.POINTER(rtc_stats)]
    if hasattr(L, "rtc_static_max_tiles"):  # (absent in older builds that tools/kernel_variants.py compares against)
        L.rtc_static_max_tiles.restype = ctypes.c_uint64
        L.rtc_render_device_static.argtypes = [vp, ctypes.POINTER(rtc_camera), ctypes.c_int, vp, vp, ctypes.POINTER(rtc_stats)]
    L.rtc_partition_rows.argtypes = [ctypes.c_uint64, ctypes.POINTER(rtc_partition)]
    L.rtc_partition_rows.restype = ctypes.c_uint64
    L.rtc_load_scene_json.argtypes = [cp, cp, ctypes.c_long, ctypes.c_long, ctypes.POINTER(vp),
                                      ctypes.POINTER(ctypes.c_size_t), ctypes.POINTER(rtc_camera)]
    L.rtc_blob_free.argtypes = [vp]
    L.rtc_blob_free.restype = None
    L.rtc_host_alloc.argtypes = [ctypes.c_size_t]
    L.rtc_host_alloc.restype = vp
    L.rtc_host_free.argtypes = [vp]
    L.rtc_host_free.restype = None
    L.rtc_camera_look_at.argtypes = [ctypes.POINTER(rtc_camera)] + [ctypes.POINTER(ctypes.c_double)] * 3
    L.rtc_canvas_ppm.argtypes = [vp, ctypes.c_uint64, ctypes.c_uint64, ctypes.POINTER(vp), ctypes.POINTER(ctypes.c_size_t)]
    L.rtc_write_png.argtypes = [cp, vp, ctypes.c_uint64, ctypes.c_uint64]
    L.startInitRenderer.argtypes = [cp, ctypes.c_size_t]
    L.startInitRenderer.restype = None
    for name in ("tryFinishInitRenderer", "initRendererIsOk", "tryFinishRender"):
        getattr(L, name).restype = ctypes.c_bool
    L.initRendererGetPixels.restype = vp
    L.initRendererGetWidth.restype = ctypes.c_size_t
    L.initRendererGetHeight.restype = ctypes.c_size_t
    L.initRendererGetErr.restype = cp
    for name in ("deinitRenderer", "startRender"):
        getattr(L, name).restype = None
    for name in ("rotateCamera", "moveCamera"):
        getattr(L, name).argtypes = [ctypes.c_double]
        getattr(L, name).restype = None
    L.rtcRendererSetAssetDir.argtypes = [cp]
    L.rtcRendererSetAssetDir.restype = None
    L.rtcRendererLastStats.argtypes = [ctypes.POINTER(rtc_stats)]
    _lib = L
    return L


def _finish_old(L):
    """Bindings of a round-1 build: the entry points both rounds have (enough for tools/bench_all.py)."""
    global _lib
    vp, cp = ctypes.c_void_p, ctypes.c_char_p
    L.rtc_scene_create.argtypes = [vp, ctypes.c_size_t, ctypes.POINTER(vp)]
    L.rtc_scene_destroy.argtypes = [vp]
    L.rtc_scene_destroy.restype = None
    common = [vp, ctypes.POINTER(rtc_camera), ctypes.c_int, ctypes.POINTER(rtc_partition)]
    L.rtc_render_device.argtypes = common + [vp, vp, ctypes.POINTER(rtc_stats)]
    L.rtc_render_device_frame.argtypes = common + [vp, vp, ctypes.POINTER(rtc_stats)]
    L.rtc_load_scene_json.argtypes = [cp, cp, ctypes.c_long, ctypes.c_long, ctypes.POINTER(vp),
                                      ctypes.POINTER(ctypes.c_size_t), ctypes.POINTER(rtc_camera)]
    L.rtc_blob_free.argtypes = [vp]
    L.rtc_blob_free.restype = None
    L.rtc_set_devices = lambda n: 0
    _lib = L
    return L


def _check(rc):
    if rc != 0:
        raise RtcError(rc, lib().rtc_last_error().decode(errors="replace"))


def device_count():
    return lib().rtc_device_count()


def load_blob(json_text, asset_dir="", width=0, height=0):
    """Host loader (C++ mirror of src/parsing): -> (blob bytes, rtc_camera).  Needs no GPU."""
    L = lib()
    blob, n, cam = ctypes.c_void_p(), ctypes.c_size_t(), rtc_camera()
    _check(L.rtc_load_scene_json(json_text.encode(), asset_dir.encode(), width, height, ctypes.byref(blob),
                                 ctypes.byref(n), ctypes.byref(cam)))
    try:
        data = ctypes.string_at(blob, n.value)
    finally:
        L.rtc_blob_free(blob)
    return data, cam


def inspect_blob(blob):
    """What rtc_scene_create would build from this blob (prims, mesh instances, shared geometries, BVH size): host only."""
    info = rtc_scene_info()
    _check(lib().rtc_blob_inspect(blob, len(blob), ctypes.byref(info)))
    return info


class _Pinned:
    def __init__(self, ptr):
        self.ptr = ptr

    def __del__(self):
        try:
            lib().rtc_host_free(self.ptr)
        except Exception:
            pass


def pinned_empty(shape, dtype=np.float64):
    """numpy array in page-locked, device-mapped host memory (rtc_host_alloc).  Camera.render(out=...) into such an
    array has the kernel write the canvas directly, with no device->host copy after the render."""
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    p = lib().rtc_host_alloc(n)
    if not p:
        raise RtcError(RTC_ERR_CUDA, lib().rtc_last_error().decode(errors="replace"))
    owner = _Pinned(p)
    buf = (ctypes.c_char * n).from_address(p)
    arr = np.frombuffer(buf, dtype=dtype).reshape(shape)
    _PINNED_OWNERS[arr.ctypes.data] = owner  # freed with free_pinned(); numpy cannot own foreign memory
    return arr


_PINNED_OWNERS = {}


def free_pinned(arr):
    _PINNED_OWNERS.pop(arr.ctypes.data, None)


class Canvas:
    """canvas.zig:18-21: width, height, pixels (linear RGB)."""

    def __init__(self, pixels, stats=None):
        self.pixels = pixels
        self.height, self.width = pixels.shape[:2]
        self.stats = stats

    def to_image(self):
        """Canvas.toImage (canvas.zig:150-176) with color.zig:61-71 clamp(): round-half-away(c*255) in [0,255]."""
        t = self.pixels * 255.0
        r = np.where(t >= 0, np.floor(t + 0.5), -np.floor(-t + 0.5))
        return np.clip(np.nan_to_num(r, nan=0.0), 0, 255).astype(np.uint8)

    def ppm(self):
        """Canvas.ppm (canvas.zig:181-254), character for character (lines broken before column 70)."""
        L = lib()
        px = np.ascontiguousarray(self.pixels, dtype=np.float64)
        text, n = ctypes.c_void_p(), ctypes.c_size_t()
        _check(L.rtc_canvas_ppm(px.ctypes.data, self.width, self.height, ctypes.byref(text), ctypes.byref(n)))
        try:
            return ctypes.string_at(text, n.value).decode()
        finally:
            L.rtc_blob_free(text)

    def save_png(self, path):
        """main.zig:95-99: canvas.toImage() written as an rgb24 PNG."""
        img = self.to_image()
        rgba = np.concatenate([img, np.full(img.shape[:2] + (1,), 255, dtype=np.uint8)], axis=2)
        _check(lib().rtc_write_png(str(path).encode(), np.ascontiguousarray(rgba).ctypes.data, self.width, self.height))


class World:
    """Device-resident world (world.zig:20-25 after flattening).  Owns the rtc_scene handle.
    devices > 1: the scene is replicated on `devices` GPUs of this process (rtc_set_devices) and whole-frame renders use
    all of them."""

    def __init__(self, blob, device=0, devices=1):
        L = lib()
        if L.rtc_device_count() <= 0:
            raise RtcError(RTC_ERR_CUDA, "no CUDA device visible; this renderer has no CPU fallback")
        _check(L.rtc_set_device(device))
        _check(L.rtc_set_devices(devices))
        self._h = ctypes.c_void_p()
        self._blob_bytes = len(blob)
        try:
            _check(L.rtc_scene_create(blob, len(blob), ctypes.byref(self._h)))
        finally:
            L.rtc_set_devices(1)

    def share_tiles(self, rank, n_ranks, words, cost_ptrs):
        """rtc_scene_share_tiles: cost_ptrs[b][r] = device address (int) of buffer b of renderer r.  n_ranks <= 1 ends it."""
        if n_ranks <= 1:
            _check(lib().rtc_scene_share_tiles(self._h, None))
            return
        sh = rtc_tile_share()
        sh.rank, sh.n_ranks, sh.words = rank, n_ranks, words
        for b in range(2):
            for r in range(n_ranks):
                sh.cost[b][r] = cost_ptrs[b][r]
        _check(lib().rtc_scene_share_tiles(self._h, ctypes.byref(sh)))

    def fetch_stats(self, max_frames=16):
        """Waits for the frames put on a stream by Camera.render_device_async and returns their rtc_stats, oldest first."""
        arr = (rtc_stats * max_frames)()
        n = ctypes.c_uint32()
        _check(lib().rtc_stats_fetch(self._h, arr, max_frames, ctypes.byref(n)))
        return [arr[i] for i in range(n.value)]

    def close(self):
        if getattr(self, "_h", None):
            lib().rtc_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Camera:
    """camera.zig:16-125.  render() is the drop-in for Camera.render."""

    def __init__(self, cam):
        self.c = cam

    @property
    def hsize(self):
        return int(self.c.hsize)

    @property
    def vsize(self):
        return int(self.c.vsize)

    def look_at(self, frm, to, up):
        """lib.zig:166-190 (rotateCamera / moveCamera recompute the view from from/to/up)."""
        a = [(ctypes.c_double * 3)(*v) for v in (frm, to, up)]
        _check(lib().rtc_camera_look_at(ctypes.byref(self.c), *a))

    @staticmethod
    def _part(partition):
        if partition is None:
            return None, None
        p = rtc_partition(*partition)
        return p, ctypes.byref(p)

    def render(self, world, max_depth=5, partition=None, out=None):
        """-> Canvas.  partition = (tile_rows, part, n_parts) renders only those rows (others stay 0)."""
        h, w = self.vsize, self.hsize
        if out is None:
            out = np.zeros((h, w, 3), dtype=np.float64)
        assert out.dtype == np.float64 and out.shape == (h, w, 3) and out.flags["C_CONTIGUOUS"]
        st = rtc_stats()
        keep, pref = self._part(partition)
        _check(lib().rtc_render(world._h, ctypes.byref(self.c), max_depth, pref, out.ctypes.data, ctypes.byref(st)))
        return Canvas(out, st)

    def render_rgba8(self, world, max_depth=5, partition=None):
        """lib.zig:135-155 twin -> ([H,W,4] uint8, stats)."""
        h, w = self.vsize, self.hsize
        out = np.zeros((h, w, 4), dtype=np.uint8)
        st = rtc_stats()
        keep, pref = self._part(partition)
        _check(lib().rtc_render_rgba8(world._h, ctypes.byref(self.c), max_depth, pref, out.ctypes.data,
                                      ctypes.byref(st)))
        return out, st

    def partition_rows(self, partition):
        keep, pref = self._part(partition)
        return int(lib().rtc_partition_rows(self.vsize, pref))

    def render_device(self, world, d_ptr, max_depth=5, partition=None, stream=0):
        """Result stays in device memory at d_ptr (packed partition rows, float64 RGB)."""
        st = rtc_stats()
        keep, pref = self._part(partition)
        _check(lib().rtc_render_device(world._h, ctypes.byref(self.c), max_depth, pref, ctypes.c_void_p(d_ptr),
                                       ctypes.c_void_p(stream), ctypes.byref(st)))
        return st


    def render_device_frame(self, world, d_frame_ptr, max_depth=5, partition=None, stream=0):
        """Rows of `partition` written at their positions of a FULL frame at d_frame_ptr (float64 RGB) -- possibly
        peer-mapped memory of another GPU (multi_gpu.PeerFrame)."""
        st = rtc_stats()
        keep, pref = self._part(partition)
        _check(lib().rtc_render_device_frame(world._h, ctypes.byref(self.c), max_depth, pref,
                                             ctypes.c_void_p(d_frame_ptr), ctypes.c_void_p(stream), ctypes.byref(st)))
        return st


    def render_device_static(self, world, d_ptr, max_depth=5, stream=0, wait=True):
        """rtc_render_device_static: the whole frame into device memory at d_ptr WITHOUT any allocation (scratch in a
        fixed array of the library's device image; frames of at most rtc_static_max_tiles() 8x4-pixel tiles).
        wait=False only puts the work on `stream` and returns None."""
        st = rtc_stats() if wait else None
        _check(lib().rtc_render_device_static(world._h, ctypes.byref(self.c), max_depth, ctypes.c_void_p(d_ptr),
                                              ctypes.c_void_p(stream), ctypes.byref(st) if wait else None))
        return st

    def render_device_async(self, world, d_frame_ptr, max_depth=5, partition=None, stream=0):
        """render_device_frame without the wait: the frame is put on `stream`; World.fetch_stats() collects the statistics."""
        keep, pref = self._part(partition)
        _check(lib().rtc_render_device_async(world._h, ctypes.byref(self.c), max_depth, pref,
                                             ctypes.c_void_p(d_frame_ptr), ctypes.c_void_p(stream)))


class SceneInfo:
    """scene.zig:608-610: { camera, world }."""

    def __init__(self, camera, world):
        self.camera, self.world = camera, world


def parse_scene(json_text, asset_dir="", width=0, height=0, device=0):
    """parseScene (scene.zig:612-661) + upload.  width/height override camera.width/height."""
    blob, cam = load_blob(json_text, asset_dir, width, height)
    return SceneInfo(Camera(cam), World(blob, device))


def parse_scene_file(path, asset_dir="", width=0, height=0, device=0):
    with open(path) as f:
        return parse_scene(f.read(), asset_dir, width, height, device)


class Renderer:
    """The reference's global Renderer (lib.zig:65-191, 194) driven the way www/index.js drives it: start, poll,
    read the library-owned RGBA8 buffer, orbit, render again.  One per process."""

    def __init__(self, json_text, asset_dir=".", n_threads=0, poll_s=0.0005):
        import time
        L = lib()
        self._poll = poll_s
        L.rtcRendererSetAssetDir(asset_dir.encode())
        self._text = json_text.encode()  # must outlive the init worker (lib.zig:233: the pointer is read there)
        L.startInitRenderer(self._text, n_threads)
        while not L.tryFinishInitRenderer():
            time.sleep(poll_s)
        if not L.initRendererIsOk():
            raise RtcError(RTC_ERR_PARSE, L.initRendererGetErr().decode(errors="replace"))
        self.width, self.height = int(L.initRendererGetWidth()), int(L.initRendererGetHeight())
        buf = (ctypes.c_uint8 * (4 * self.width * self.height)).from_address(L.initRendererGetPixels())
        self.pixels = np.frombuffer(buf, dtype=np.uint8).reshape(self.height, self.width, 4)  # the library's buffer

    def render(self):
        """startRender + tryFinishRender until done -> the (library-owned) [H,W,4] uint8 view."""
        import time
        L = lib()
        L.startRender()
        while not L.tryFinishRender():
            time.sleep(self._poll)
        return self.pixels

    def rotate_camera(self, angle):
        lib().rotateCamera(angle)

    def move_camera(self, distance):
        lib().moveCamera(distance)

    def stats(self):
        st = rtc_stats()
        return st if lib().rtcRendererLastStats(ctypes.byref(st)) == 0 else None

    def close(self):
        self.pixels = None
        lib().deinitRenderer()
