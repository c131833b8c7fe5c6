"""Seeded inputs for the State::update parity tests (shared by tests/golden/make_golden.py and tests/test_gpu_temporal.py)."""
import numpy as np

CAMS = [((0.0, 50.0, 1000.0), (0.0, 50.0, 0.0)), ((30.0, 60.0, 990.0), (5.0, 45.0, 0.0)), ((31.0, 60.5, 989.0), (5.0, 45.0, 0.0)), ((31.0, 60.5, 989.0), (5.0, 45.0, 0.0)),
        ((-200.0, 80.0, 900.0), (0.0, 0.0, 0.0))]


def frame_inputs(po, W, H, frame, seed=0, fov=60.0, aspect=16 / 9):
    """data (rgb, 1), position (xyz, t), packed ids and the world->clip matrix of frame `frame` of a camera path over a noisy depth
    field.  Ids come in 6x5-texel patches so that some reprojections validate and some do not; a few texels carry NaN / inf / huge values."""
    rng = np.random.default_rng(seed * 1000 + frame)
    cam = CAMS[frame % len(CAMS)]
    ys, xs = np.mgrid[0:H, 0:W]
    st = np.stack([(xs + 0.5) / W, (ys + 0.5) / H], -1).reshape(-1, 2).astype(np.float32)
    o, d = po.camera_rays(cam[0], cam[1], fov, aspect, st)
    depth = (700.0 + 500.0 * rng.random(W * H)).astype(np.float32)
    pos = np.concatenate([o + d * depth[:, None], depth[:, None]], 1).reshape(H, W, 4).astype(np.float32)
    data = rng.random((H, W, 4), np.float32) * np.float32(2.0)
    data[..., 3] = 1.0
    flat = (ys // 5) * 7 + (xs // 6)
    cur = (flat % 5).astype(np.uint32)
    prev = ((flat + (rng.random((H, W)) < 0.2)) % 5).astype(np.uint32)
    ids = (prev << np.uint32(16)) | cur
    if W * H >= 64:
        k = rng.integers(0, W * H, 6)
        data.reshape(-1, 4)[k[0], 0] = np.nan
        data.reshape(-1, 4)[k[1], 1] = np.inf
        data.reshape(-1, 4)[k[2], 2] = 1e30
        pos.reshape(-1, 4)[k[3], :3] = 1e30
        pos.reshape(-1, 4)[k[4], 0] = np.nan
        pos.reshape(-1, 4)[k[5], :3] = (cam[0][0], cam[0][1], cam[0][2] + 10.0)  # behind the camera: w < 1
    vp = po.view_projection(cam[0], cam[1], fov, aspect).astype(np.float32)
    return data, pos, ids, vp


def run_sequence(update, read, po, W, H, frames, seed=0):
    """Drives `update(data, pos, ids, inv, last)` through `frames` frames the way main.rs:208-216 does; returns read() per frame."""
    last = frame_inputs(po, W, H, 0, seed)[3]  # main.rs:128: the first frame sees an unchanged camera
    out = []
    for f in range(frames):
        data, pos, ids, vp = frame_inputs(po, W, H, f, seed)
        update(data, pos, ids, vp, last)
        last = vp
        out.append(read())
    return out
