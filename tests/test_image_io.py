"""Output / ingest stages either side of the path: PNG coding, env-map ingest, and the oracle's
restatement of write_image's tonemap (known points derived from tonemapping.rs by hand)."""
import zlib
import struct

import numpy as np

from path_tracer_b200 import image_io


def test_png_roundtrip_and_env_gamma(tmp_path):
    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, (17, 23, 3), dtype=np.uint8)
    p = tmp_path / "a.png"
    image_io.write_png(p, img)
    back = image_io.read_png(p)
    assert back.dtype == np.float32 and back.shape == (17, 23, 3)
    assert np.array_equal(back, img.astype(np.float32) / np.float32(255.0))   # into_rgb32f: u8 / 255
    env = image_io.load_env(p)
    assert np.allclose(env, np.power(img.astype(np.float64) / 255.0, 2.2), rtol=2e-6, atol=1e-7)  # powf(2.2), image_helper.rs:30


def test_png_decoder_handles_all_filters(tmp_path):
    # hand-built PNG using filter types 1..4 on successive rows (an encoder other than ours)
    h, w = 5, 7
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    flat = img.reshape(h, w * 3).astype(np.int32)
    bpp, rows, prev = 3, [], np.zeros(w * 3, np.int32)
    for y in range(h):
        ft = y % 5
        cur, out = flat[y], np.zeros(w * 3, np.int32)
        for i in range(w * 3):
            a = cur[i - bpp] if i >= bpp else 0
            b = prev[i]
            c = prev[i - bpp] if i >= bpp else 0
            if ft == 0:
                pr = 0
            elif ft == 1:
                pr = a
            elif ft == 2:
                pr = b
            elif ft == 3:
                pr = (a + b) >> 1
            else:
                pa, pb, pc = abs(b - c), abs(a - c), abs(a + b - 2 * c)
                pr = a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)
            out[i] = (cur[i] - pr) & 0xFF
        rows.append(bytes([ft]) + out.astype(np.uint8).tobytes())
        prev = cur
    ch = lambda t, d: struct.pack(">I", len(d)) + t + d + struct.pack(">I", zlib.crc32(t + d) & 0xFFFFFFFF)
    png = b"\x89PNG\r\n\x1a\n" + ch(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)) + ch(b"IDAT", zlib.compress(b"".join(rows))) + ch(b"IEND", b"")
    p = tmp_path / "f.png"
    p.write_bytes(png)
    assert np.array_equal(image_io.read_png(p), img.astype(np.float32) / np.float32(255.0))


def test_oracle_tonemap_known_points(po):
    # count w = 2: the display value is rgb / w (shader.wgsl:63)
    xs = np.float32([-1.0, 0.0, 0.22, 0.44, 0.53, 1.0, 4.0, 200.0, np.nan])
    a = np.zeros((1, len(xs), 4), np.float32)
    a[..., 3] = 2
    a[0, :, 0] = a[0, :, 1] = a[0, :, 2] = xs * 2
    out = po.tonemap_rgb8(a, flip_y=False)[0, :, 0].astype(int)
    m, l0 = 0.22, (1 - 0.22) * 0.4
    lin = lambda x: int((x ** (1 / 2.2)) * 255.0)          # linear section: gt_linear with a = 1 is the identity
    assert out[0] == 0 and out[1] == 0 and out[8] == 0      # x < 0 -> b = 0; 0; NaN `as u8` = 0
    assert abs(out[2] - lin(0.22)) <= 1 and abs(out[3] - lin(0.44)) <= 1 and abs(out[4] - lin(0.53)) <= 1
    sh = lambda x: 1 - (1 - (m + l0)) * np.exp(-(1.0 / (1 - (m + l0))) * (x - (m + l0)))   # gt_shoulder with p = a = 1
    assert abs(out[5] - int(sh(1.0) ** (1 / 2.2) * 255.0)) <= 1 and out[6] >= 250 and out[7] == 255
    assert np.all(np.diff(out[1:8]) >= 0)
    # exactly at x = m + l0 the shoulder weight gt_lerp(x, e, e) is 0/0 = NaN upstream (tonemapping.rs:84, :21-35): NaN `as u8` = 0
    e = np.float32(np.float32(np.float32(1.0) - np.float32(0.22)) * np.float32(0.4)) + np.float32(0.22)
    q = np.zeros((1, 1, 4), np.float32)
    q[0, 0] = (e, e, e, 1)
    assert po.tonemap_rgb8(q)[0, 0, 0] == 0
    # flip_y puts the bottom accumulation row last -> first
    b = np.zeros((2, 1, 4), np.float32)
    b[..., 3] = 1
    b[1, 0, :3] = 1.0
    f = po.tonemap_rgb8(b, flip_y=True)
    assert f[0, 0, 0] > 200 and f[1, 0, 0] == 0
