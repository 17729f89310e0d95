"""The host loader's PNG decoder (csrc/host/images.cpp) must reject crafted files instead of indexing past its buffers
(ADVICE r1): IHDR must come first and be 13 bytes, sizes must be positive and bounded."""
import json
import struct
import zlib

import pytest


def chunk(t, d):
    return struct.pack(">I", len(d)) + t + d + struct.pack(">I", zlib.crc32(t + d) & 0xffffffff)


SIG = b"\x89PNG\r\n\x1a\n"


def ihdr(w, h, depth=8, ctype=2, interlace=0):
    return chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, ctype, 0, 0, interlace))


FILES = {
    "good.png": (SIG + ihdr(2, 2) + chunk(b"IDAT", zlib.compress(b"\x00" + b"\x10" * 6 + b"\x00" + b"\x20" * 6)) + chunk(b"IEND", b""), True),
    "idat_first.png": (SIG + chunk(b"IDAT", zlib.compress(b"")) + chunk(b"IEND", b""), False),
    "short_ihdr.png": (SIG + chunk(b"IHDR", b"\0" * 8) + chunk(b"IEND", b""), False),
    "huge.png": (SIG + ihdr(0xffffffff, 0xffffffff) + chunk(b"IDAT", zlib.compress(b"\0")) + chunk(b"IEND", b""), False),
    "zero_width.png": (SIG + ihdr(0, 5) + chunk(b"IDAT", zlib.compress(b"\0")) + chunk(b"IEND", b""), False),
    "sixteen_bit.png": (SIG + ihdr(2, 2, depth=16) + chunk(b"IDAT", zlib.compress(b"\0" * 26)) + chunk(b"IEND", b""), False),
    "interlaced.png": (SIG + ihdr(2, 2, interlace=1) + chunk(b"IDAT", zlib.compress(b"\0" * 14)) + chunk(b"IEND", b""), False),
}


def scene(file_name):
    return json.dumps({
        "camera": {"width": 8, "height": 8, "field-of-view": 1.0, "from": [0, 0, -5], "to": [0, 0, 0], "up": [0, 1, 0]},
        "lights": [{"point-light": {"position": [0, 5, -5], "intensity": [1, 1, 1]}}],
        "objects": [{"type": {"sphere": {}}, "material": {"pattern": {"type": {"texture-map": {"spherical": {
            "uv-pattern": {"image": {"file": file_name}}}}}}}}]})


@pytest.mark.parametrize("name", sorted(FILES))
def test_png_files(pkg, tmp_path, name):
    data, ok = FILES[name]
    (tmp_path / name).write_bytes(data)
    if ok:
        blob, _ = pkg.load_blob(scene(name), str(tmp_path))
        assert len(blob) > 0
    else:
        with pytest.raises(pkg.RtcError):
            pkg.load_blob(scene(name), str(tmp_path))
