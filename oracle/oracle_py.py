"""ORACLE — TEST INFRASTRUCTURE ONLY.  ctypes wrapper around oracle/liboracle.so (the C++ restatement
of the reference's CPU render path, see oracle/oracle.hpp).  Importable only from tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    if force or not os.path.exists(so):
        subprocess.check_call(["make", "-C", _HERE, "all"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = ctypes.CDLL(build())
        L.oracle_scene_load.restype = ctypes.c_void_p
        L.oracle_scene_load.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_long, ctypes.c_long,
                                        ctypes.c_char_p, ctypes.c_size_t]
        L.oracle_scene_parse.restype = ctypes.c_void_p
        L.oracle_scene_parse.argtypes = L.oracle_scene_load.argtypes
        L.oracle_scene_dims.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_long), ctypes.POINTER(ctypes.c_long)]
        L.oracle_render.restype = ctypes.c_int
        L.oracle_render.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_long, ctypes.c_long, ctypes.c_int,
                                    ctypes.c_void_p, ctypes.POINTER(ctypes.c_ulonglong)]
        L.oracle_set_camera.argtypes = [ctypes.c_void_p] + [ctypes.POINTER(ctypes.c_double)] * 3
        L.oracle_scene_free.argtypes = [ctypes.c_void_p]
        _LIB = L
    return _LIB


class OracleScene:
    def __init__(self, json_path=None, asset_dir="", width=0, height=0, json_text=None):
        err = ctypes.create_string_buffer(512)
        if json_text is not None:
            h = lib().oracle_scene_parse(json_text.encode(), asset_dir.encode(), width, height, err, 512)
        else:
            h = lib().oracle_scene_load(json_path.encode(), asset_dir.encode(), width, height, err, 512)
        if not h:
            raise RuntimeError("oracle: " + err.value.decode())
        self._h = h
        w, hh = ctypes.c_long(), ctypes.c_long()
        lib().oracle_scene_dims(h, ctypes.byref(w), ctypes.byref(hh))
        self.width, self.height = w.value, hh.value

    def render(self, threads=None, y0=0, y1=-1, max_depth=5):
        """-> (float64 [H,W,3] linear RGB, dict ray counts)."""
        threads = threads or os.cpu_count() or 1
        out = np.zeros((self.height, self.width, 3), dtype=np.float64)
        counts = (ctypes.c_ulonglong * 3)()
        rc = lib().oracle_render(self._h, threads, y0, y1, max_depth, out.ctypes.data, counts)
        if rc != 0:
            raise RuntimeError("oracle render failed")
        return out, {"primary": counts[0], "shadow": counts[1], "secondary": counts[2]}

    def set_camera(self, frm, to, up):
        a = [(ctypes.c_double * 3)(*v) for v in (frm, to, up)]
        lib().oracle_set_camera(self._h, *a)

    def close(self):
        if self._h:
            lib().oracle_scene_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def quantise(rgb):
    """color.zig:61-71 clamp(): round-half-away(c*255) clipped to [0,255]."""
    t = rgb * 255.0
    r = np.where(t >= 0, np.floor(t + 0.5), -np.floor(-t + 0.5))
    return np.clip(np.nan_to_num(r, nan=0.0), 0, 255).astype(np.uint8)
