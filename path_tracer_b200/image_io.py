"""Image files either side of the render path, host-side data-format code as upstream:

* `write_png`  — the 8-bit image ptb_read_rgb8 produces (GT tonemap + gamma, image_helper.rs:37-58) as an RGB PNG.
* `load_env`   — ImageHelper::load_image (image_helper.rs:25-33): decode an image, `into_rgb32f` (u8 / 255, u16 / 65535),
                 gamma 2.2 -> linear with powf(2.2); the result feeds ptb_scene_set_env / the lat-long lookup of
                 integrator.rs:254-262.
PNG coding is done here with zlib only (8/16-bit grey, grey+alpha, RGB, RGBA, palette; non-interlaced) so that neither the
product nor its tests depend on an imaging library.
"""
from __future__ import annotations

import struct
import zlib

import numpy as np

_SIG = b"\x89PNG\r\n\x1a\n"


def _chunk(tag: bytes, data: bytes) -> bytes:
    return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)


def write_png(path, rgb8: np.ndarray, level: int = 6) -> None:
    """rgb8: [H, W, 3] uint8, first row = top of the image."""
    a = np.ascontiguousarray(rgb8, np.uint8)
    if a.ndim != 3 or a.shape[2] != 3:
        raise ValueError("expected an [H, W, 3] uint8 array")
    h, w = a.shape[:2]
    raw = np.zeros((h, 1 + w * 3), np.uint8)  # filter type 0 on every scanline
    raw[:, 1:] = a.reshape(h, w * 3)
    data = _SIG + _chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)) + _chunk(b"IDAT", zlib.compress(raw.tobytes(), level)) + _chunk(b"IEND", b"")
    with open(path, "wb") as f:
        f.write(data)


def _unfilter(raw: np.ndarray, h: int, stride: int, bpp: int) -> np.ndarray:
    out = np.zeros((h, stride), np.uint8)
    prev = np.zeros(stride, np.int32)
    pos = 0
    for y in range(h):
        ft = int(raw[pos])
        line = raw[pos + 1: pos + 1 + stride].astype(np.int32)
        pos += 1 + stride
        if ft == 0:
            cur = line
        elif ft == 2:
            cur = (line + prev) & 0xFF
        elif ft in (1, 3, 4):
            cur = np.zeros(stride, np.int32)
            for i in range(stride):  # left-dependent filters: sequential per byte
                a = cur[i - bpp] if i >= bpp else 0
                b = prev[i]
                if ft == 1:
                    p = a
                elif ft == 3:
                    p = (a + b) >> 1
                else:
                    c = prev[i - bpp] if i >= bpp else 0
                    pa, pb, pc = abs(b - c), abs(a - c), abs(a + b - 2 * c)
                    p = a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)
                cur[i] = (line[i] + p) & 0xFF
        else:
            raise ValueError(f"bad PNG filter type {ft}")
        out[y] = cur
        prev = cur
    return out


def read_png(path) -> np.ndarray:
    """Decode a non-interlaced PNG to float32 RGB in [0, 1] as image::DynamicImage::into_rgb32f does (sample / max)."""
    with open(path, "rb") as f:
        data = f.read()
    if data[:8] != _SIG:
        raise ValueError("not a PNG file")
    pos, idat, plte, ihdr = 8, [], None, None
    while pos < len(data):
        n, tag = struct.unpack(">I4s", data[pos: pos + 8])
        body = data[pos + 8: pos + 8 + n]
        pos += 12 + n
        if tag == b"IHDR":
            ihdr = struct.unpack(">IIBBBBB", body)
        elif tag == b"PLTE":
            plte = np.frombuffer(body, np.uint8).reshape(-1, 3)
        elif tag == b"IDAT":
            idat.append(body)
        elif tag == b"IEND":
            break
    if ihdr is None:
        raise ValueError("PNG without IHDR")
    w, h, depth, ctype, _, _, interlace = ihdr
    if interlace:
        raise ValueError("interlaced PNG not supported")
    if depth not in (8, 16) or ctype not in (0, 2, 3, 4, 6) or (ctype == 3 and depth != 8):
        raise ValueError(f"unsupported PNG format (depth {depth}, colour type {ctype})")
    channels = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ctype]
    bpp = channels * depth // 8
    stride = w * bpp
    raw = np.frombuffer(zlib.decompress(b"".join(idat)), np.uint8)
    px = _unfilter(raw, h, stride, bpp)
    if depth == 16:
        v = px.reshape(h, w, channels, 2).astype(np.uint32)
        s = (v[..., 0] << 8 | v[..., 1]).astype(np.float32) / np.float32(65535.0)
    elif ctype == 3:
        if plte is None:
            raise ValueError("palette PNG without PLTE")
        s = plte[px.reshape(h, w)].astype(np.float32) / np.float32(255.0)
        channels = 3
    else:
        s = px.reshape(h, w, channels).astype(np.float32) / np.float32(255.0)
    if channels in (1, 2):
        return np.repeat(s[..., :1], 3, axis=2)
    return np.ascontiguousarray(s[..., :3])


def load_env(path) -> np.ndarray:
    """ImageHelper::load_image (image_helper.rs:25-33): [H, W, 3] float32, linear RGB = decoded ^ 2.2; row 0 = first image row
    (the lookup of integrator.rs:256-262 / image_helper.rs:71-88 indexes it directly)."""
    return np.power(read_png(path), np.float32(2.2)).astype(np.float32)
