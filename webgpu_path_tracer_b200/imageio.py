"""Headless stand-in for the canvas / swap chain of the reference.

The reference presents `fs`'s output on a WebGPU canvas (/root/reference/renderer.js:130-136,
shaders/fragment.js:22-36); there is no window system here, so the presented RGBA8 image goes to a
file instead: binary PPM (P6) or PNG (8-bit RGB / RGBA, zlib from the standard library, no
third-party imaging dependency)."""
import struct
import zlib

import numpy as np


def to_rgba8(image):
    """float RGBA in [0,1] (what wgpt_present returns) or uint8 -> uint8 (H, W, 4), round(c * 255)
    exactly as wgpt_present's own rgba8 output."""
    a = np.asarray(image)
    if a.dtype == np.uint8:
        return a
    return np.clip(np.rint(np.clip(a, 0.0, 1.0) * 255.0), 0, 255).astype(np.uint8)


def write_ppm(path, image):
    a = to_rgba8(image)
    h, w = a.shape[:2]
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (w, h))
        f.write(np.ascontiguousarray(a[:, :, :3]).tobytes())


def _chunk(tag, data):
    body = tag + data
    return struct.pack(">I", len(data)) + body + struct.pack(">I", zlib.crc32(body) & 0xffffffff)


def encode_png(image, alpha=False):
    a = to_rgba8(image)
    h, w = a.shape[:2]
    px = np.ascontiguousarray(a if alpha else a[:, :, :3])
    rows = np.concatenate([np.zeros((h, 1), np.uint8), px.reshape(h, -1)], axis=1)      # filter type 0 per scanline
    ihdr = struct.pack(">IIBBBBB", w, h, 8, 6 if alpha else 2, 0, 0, 0)
    return b"\x89PNG\r\n\x1a\n" + _chunk(b"IHDR", ihdr) + _chunk(b"IDAT", zlib.compress(rows.tobytes(), 6)) + _chunk(b"IEND", b"")


def write_png(path, image, alpha=False):
    with open(path, "wb") as f:
        f.write(encode_png(image, alpha))


def decode_png(data):
    """Minimal reader for the files encode_png writes (tests round-trip through it)."""
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w = 8, b"", None
    while pos < len(data):
        n, = struct.unpack(">I", data[pos:pos + 4])
        tag, body = data[pos + 4:pos + 8], data[pos + 8:pos + 8 + n]
        crc, = struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])
        assert crc == (zlib.crc32(tag + body) & 0xffffffff)
        if tag == b"IHDR":
            w, h, depth, ctype = struct.unpack(">IIBB", body[:10])
            assert depth == 8 and ctype in (2, 6)
            ch = 4 if ctype == 6 else 3
        elif tag == b"IDAT":
            idat += body
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + w * ch)
    assert not raw[:, 0].any()
    return raw[:, 1:].reshape(h, w, ch)


def write_pfm(path, image):
    """HDR output of the f32 present (or of binding 3 itself): Portable Float Map, little-endian RGB float32, rows
    bottom-to-top (the format's convention).  `image`: (H, W, 3) or (H, W, 4) floats."""
    a = np.asarray(image, np.float32)
    h, w = a.shape[:2]
    with open(path, "wb") as f:
        f.write(b"PF\n%d %d\n-1.0\n" % (w, h))
        f.write(np.ascontiguousarray(a[::-1, :, :3]).astype("<f4").tobytes())


def read_pfm(path):
    with open(path, "rb") as f:
        assert f.readline().strip() == b"PF"
        w, h = (int(x) for x in f.readline().split())
        scale = float(f.readline())
        data = np.frombuffer(f.read(), "<f4" if scale < 0 else ">f4").reshape(h, w, 3)
    return data[::-1].astype(np.float32)
