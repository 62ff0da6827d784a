"""Headless image output for ``sail_present``.

The reference shows its display pass on a canvas (/root/reference/src/core/renderer.js:63, main/fsrender.glsl:4-6);
with no window here the RGBA8 buffer is written to a file instead.  Device images have row 0 at the BOTTOM
(gl_FragCoord); files are written top row first, as a canvas shows them.  PNG is encoded with zlib only (8-bit
RGB/RGBA, filter 0, no interlace); ``read_png`` decodes what ``write_png`` writes (used by the tests) and ordinary
non-interlaced 8-bit RGB/RGBA PNGs with any of the five row filters.
"""
import struct
import zlib

import numpy as np


def _as_u8(img):
    a = np.asarray(img)
    if a.dtype != np.uint8:
        # the display pass's float output: clamp to [0, 1] and quantise as an RGBA8 framebuffer does (round to nearest)
        a = np.nan_to_num(a.astype(np.float32), nan=0.0, posinf=1.0, neginf=0.0)
        a = np.floor(np.clip(a, 0.0, 1.0) * 255.0 + 0.5).astype(np.uint8)
    if a.ndim != 3 or a.shape[2] not in (3, 4):
        raise ValueError("expected an H x W x 3|4 image, got %r" % (a.shape,))
    return a


def to_top_down(img):
    """Bottom-row-first device image -> top-row-first file order."""
    return np.ascontiguousarray(_as_u8(img)[::-1])


def write_ppm(path, img):
    """Binary PPM (P6), RGB; alpha is dropped."""
    a = to_top_down(img)[:, :, :3]
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (a.shape[1], a.shape[0]))
        f.write(np.ascontiguousarray(a).tobytes())


def _chunk(tag, data):
    return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)


def write_png(path, img, alpha=False, level=6):
    """8-bit PNG.  ``alpha=False`` writes RGB (the canvas is opaque: fsrender.glsl writes alpha 1)."""
    a = to_top_down(img)
    a = a if (alpha and a.shape[2] == 4) else a[:, :, :3]
    h, w, c = a.shape
    raw = np.empty((h, 1 + w * c), np.uint8)
    raw[:, 0] = 0  # filter type 0 on every scanline
    raw[:, 1:] = a.reshape(h, w * c)
    ihdr = struct.pack(">IIBBBBB", w, h, 8, 6 if c == 4 else 2, 0, 0, 0)
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + _chunk(b"IHDR", ihdr) + _chunk(b"IDAT", zlib.compress(raw.tobytes(), level)) + _chunk(b"IEND", b""))


def read_png(path):
    """Decode a non-interlaced 8-bit RGB/RGBA PNG -> H x W x C uint8, top row first."""
    data = open(path, "rb").read()
    if data[:8] != b"\x89PNG\r\n\x1a\n":
        raise ValueError("not a PNG file")
    pos, idat, w = 8, b"", None
    while pos < len(data):
        n, tag = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        if zlib.crc32(tag + body) & 0xFFFFFFFF != struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])[0]:
            raise ValueError("bad CRC in %r chunk" % tag)
        if tag == b"IHDR":
            w, h, depth, ctype, _, _, interlace = struct.unpack(">IIBBBBB", body)
            if depth != 8 or ctype not in (2, 6) or interlace:
                raise ValueError("unsupported PNG flavour")
            c = 4 if ctype == 6 else 3
        elif tag == b"IDAT":
            idat += body
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + w * c)
    out = np.zeros((h, w * c), np.int32)
    prev = np.zeros(w * c, np.int32)
    for y in range(h):
        ft, line = int(raw[y, 0]), raw[y, 1:].astype(np.int32)
        if ft == 0:
            cur = line
        elif ft == 2:
            cur = (line + prev) & 255
        else:  # Sub, Average, Paeth depend on the pixel to the left: sequential over the row
            cur = np.zeros(w * c, np.int32)
            for i in range(w * c):
                a = cur[i - c] if i >= c else 0
                b = prev[i]
                cc = prev[i - c] if i >= c else 0
                if ft == 1:
                    pred = a
                elif ft == 3:
                    pred = (a + b) >> 1
                elif ft == 4:
                    p = a + b - cc
                    pa, pb, pc = abs(p - a), abs(p - b), abs(p - cc)
                    pred = a if (pa <= pb and pa <= pc) else (b if pb <= pc else cc)
                else:
                    raise ValueError("bad PNG filter type %d" % ft)
                cur[i] = (line[i] + pred) & 255
        out[y] = cur
        prev = cur
    return out.astype(np.uint8).reshape(h, w, c)
