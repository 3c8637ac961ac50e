"""Headless image output for the Film's display buffer.

The reference presents `Film::GetDisplayData()` (RGBA8, film.h:63-75) through an OpenGL texture
(main.cpp:516-532); there is no graphics stack here, so the same buffer goes to a file instead.
Writers are dependency-free: binary PPM (P6) and PNG (zlib + CRC from the standard library).
"""
import struct
import zlib

import numpy as np


def _rgb8(rgba):
    a = np.asarray(rgba)
    if a.dtype != np.uint8 or a.ndim != 3 or a.shape[2] not in (3, 4):
        raise ValueError("expected (H, W, 3|4) uint8, got %s %s" % (a.dtype, a.shape))
    return np.ascontiguousarray(a[:, :, :3])


def write_ppm(path, rgba):
    rgb = _rgb8(rgba)
    h, w = rgb.shape[:2]
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (w, h))
        f.write(rgb.tobytes())


def _chunk(tag, data):
    return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xffffffff)


def write_png(path, rgba, level=6):
    rgb = _rgb8(rgba)
    h, w = rgb.shape[:2]
    rows = np.empty((h, 1 + 3 * w), dtype=np.uint8)
    rows[:, 0] = 0                                   # filter type 0 (None) on every scanline
    rows[:, 1:] = rgb.reshape(h, 3 * w)
    png = b"\x89PNG\r\n\x1a\n"
    png += _chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0))    # 8-bit truecolour
    png += _chunk(b"IDAT", zlib.compress(rows.tobytes(), level))
    png += _chunk(b"IEND", b"")
    with open(path, "wb") as f:
        f.write(png)


def read_png_rgb8(path):
    """Inverse of write_png (filter 0 only) - used by the tests."""
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w, h = 8, b"", 0, 0
    while pos < len(data):
        n, tag = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        if tag == b"IHDR":
            w, h = struct.unpack(">II", body[:8])
        elif tag == b"IDAT":
            idat += body
        pos += 12 + n
    rows = np.frombuffer(zlib.decompress(idat), dtype=np.uint8).reshape(h, 1 + 3 * w)
    assert not rows[:, 0].any()
    return rows[:, 1:].reshape(h, w, 3).copy()


def write_image(path, rgba):
    (write_png if str(path).lower().endswith(".png") else write_ppm)(path, rgba)
