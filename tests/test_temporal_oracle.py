"""Self-checks of the oracle's State::update restatement (oracle/po_temporal.hpp; state.rs:505-579, accumulate.wgsl,
velocity.wgsl, compute.wgsl) against independent numpy arithmetic, and of the host-side view-projection of the library.
CPU only.  The reference holds no expected images for these passes (parity unpinned)."""
import numpy as np
import pytest


def _camera_inputs(po, W, H, origin, target, fov, aspect, depth, rng):
    """first-hit world positions for a camera looking at a synthetic depth field"""
    ys, xs = np.mgrid[0:H, 0:W]
    st = np.stack([(xs + 0.5) / W, (ys + 0.5) / H], -1).reshape(-1, 2).astype(np.float32)
    o, d = po.camera_rays(origin, target, fov, aspect, st)
    p = o + d * depth.reshape(-1, 1)
    return np.concatenate([p, depth.reshape(-1, 1)], 1).reshape(H, W, 4).astype(np.float32)


def test_static_camera_accumulates(po):
    rng = np.random.default_rng(1)
    H, W = 9, 13
    data = rng.random((H, W, 4), np.float32)
    data[..., 3] = 1.0
    acc0 = rng.random((H, W, 4), np.float32) * 5
    m = rng.random(16).astype(np.float32)
    acc, vel, ran = po.state_update(data, np.zeros((H, W, 4), np.float32), np.zeros((H, W), np.uint32), m, m.copy(), acc0)
    assert not ran
    want = acc0 + np.concatenate([data[..., :3], np.ones((H, W, 1), np.float32)], -1)
    assert np.array_equal(acc, want.astype(np.float32))
    assert not vel.any()


def test_velocity_matches_float64_projection(po):
    rng = np.random.default_rng(2)
    H, W = 36, 64
    cam0 = ((0.0, 50.0, 1000.0), (0.0, 50.0, 0.0))
    cam1 = ((30.0, 60.0, 990.0), (5.0, 45.0, 0.0))
    depth = (800.0 + 400.0 * rng.random((H, W))).astype(np.float32)
    pos = _camera_inputs(po, W, H, cam1[0], cam1[1], 60.0, 16 / 9, depth, rng)
    vp0 = po.view_projection(cam0[0], cam0[1], 60.0, 16 / 9)
    vp1 = po.view_projection(cam1[0], cam1[1], 60.0, 16 / 9)
    data = np.ones((H, W, 4), np.float32)
    _, vel, ran = po.state_update(data, pos, np.zeros((H, W), np.uint32), vp1.astype(np.float32), vp0.astype(np.float32), np.zeros((H, W, 4), np.float32))
    assert ran
    M = vp0.reshape(4, 4).T  # column-major -> matrix
    P = np.concatenate([pos[..., :3].astype(np.float64), np.ones((H, W, 1))], -1) @ M.T
    prev_uv = P[..., :2] / np.maximum(P[..., 3:4], 1.0) * 0.5 + 0.5
    ys, xs = np.mgrid[0:H, 0:W]
    cur_uv = np.stack([(xs + 0.5) / W, (ys + 0.5) / H], -1)
    assert np.abs(vel - (cur_uv - prev_uv)).max() < 2e-5
    # a point seen by camera 1 re-projects into camera 1 at its own pixel: zero velocity
    _, vel_same, _ = po.state_update(data, pos, np.zeros((H, W), np.uint32), vp0.astype(np.float32), vp1.astype(np.float32), np.zeros((H, W, 4), np.float32))
    assert np.abs(vel_same).max() < 1e-4  # f32 camera (tan of an f32 angle, f32 ray directions) against the exact matrix


def test_id_mismatch_gives_tent_filtered_input(po):
    """compute.wgsl:163-174: four bilinear taps at the texel corners = 3x3 tent filter with clamp-to-edge."""
    rng = np.random.default_rng(3)
    H, W = 11, 7
    data = rng.random((H, W, 4), np.float32)
    ids = np.full((H, W), 7, np.uint32)  # current id 7, previous id 0 everywhere
    a = np.eye(4, dtype=np.float32).reshape(16)
    b = a.copy()
    b[0] = 1.5  # a different matrix: the moving-camera path runs
    pos = np.zeros((H, W, 4), np.float32)
    acc, _, ran = po.state_update(data, pos, ids, a, b, np.zeros((H, W, 4), np.float32))
    assert ran
    pad = np.pad(data.astype(np.float64), ((1, 1), (1, 1), (0, 0)), mode="edge")
    k = np.array([1.0, 2.0, 1.0]) / 4.0
    want = sum(k[i] * k[j] * pad[i:i + H, j:j + W] for i in range(3) for j in range(3))
    assert np.abs(acc - want).max() < 1e-5  # uv * W - 0.5 is rounded in f32, so the weights are 0.5 only to ~1e-6


def test_reprojection_blend_properties(po):
    """Zero velocity and matching ids: output = mix(clip(history), current, 0.15) with alpha 1 (compute.wgsl:176-211)."""
    rng = np.random.default_rng(4)
    H, W = 24, 40
    cam = ((0.0, 50.0, 1000.0), (0.0, 50.0, 0.0))
    depth = np.full((H, W), 900.0, np.float32)
    pos = _camera_inputs(po, W, H, cam[0], cam[1], 60.0, 16 / 9, depth, rng)
    vp = po.view_projection(cam[0], cam[1], 60.0, 16 / 9).astype(np.float32)
    vp_now = vp.copy()
    vp_now[12] += 1e-3  # "moved" by a hair: forces the TAA path, last frame's matrix is the true one
    ids = np.full((H, W), (3 << 16) | 3, np.uint32)
    data = (0.4 + 0.2 * rng.random((H, W, 4))).astype(np.float32)
    data[..., 3] = 1.0
    # (1) history far outside the neighbourhood statistics is clipped into [mu - sigma, mu + sigma]: result stays near the input range
    hist = np.full((H, W, 4), 50.0, np.float32)
    hist[..., 3] = 1.0
    acc, _, ran = po.state_update(data, pos, ids, vp_now, vp, hist)
    assert ran and np.all(acc[..., 3] == 1.0)
    assert np.isfinite(acc).all() and acc[..., :3].max() < 1.0 and acc[..., :3].min() > 0.2
    # (2) history equal to a smooth input passes the clip almost unchanged: output close to the input
    yy, xx = np.mgrid[0:H, 0:W]
    smooth = np.stack([0.5 + 0.01 * xx, 0.3 + 0.01 * yy, 0.4 + 0.005 * (xx + yy), np.ones_like(xx)], -1).astype(np.float32)
    acc2, _, _ = po.state_update(smooth, pos, ids, vp_now, vp, smooth)
    inner = (slice(2, -2), slice(2, -2))
    assert np.abs(acc2[inner][..., :3] - smooth[inner][..., :3]).max() < 0.02
    # (3) history weight: the accumulation texture is read as rgb / max(w, 1) (w_divide): (8 rgb, 8) behaves like (rgb, 1)
    acc3, _, _ = po.state_update(smooth, pos, ids, vp_now, vp, smooth * 8.0)
    assert np.abs(acc3 - acc2).max() < 1e-5


def test_library_view_projection_matches_float64(po):
    """ptb_camera_view_projection (f32 cofactor inverse of matrix * inv_projection, as main.rs:209-216 evaluates it) against the
    exact view-projection.  Host-only context: no device involved."""
    import path_tracer_b200 as ptb
    ctx = ptb.Context(-1)
    for origin, target, fov, aspect in (((0.0, 50.0, 1000.0), (0.0, 50.0, 0.0), 60.0, 16 / 9), ((278.0, 273.0, -800.0), (278.0, 273.0, 0.0), 40.0, 1.0),
                                        ((30.0, 400.0, 990.0), (5.0, 45.0, 0.0), 75.0, 2.0)):
        ctx.set_camera(origin, target, fov, aspect)
        got = ctx.view_projection().astype(np.float64)
        want = po.view_projection(origin, target, fov, aspect)
        scale = np.abs(want).max()
        assert np.abs(got - want).max() <= 2e-5 * scale, (got, want)


def test_oracle_reproduces_temporal_golden(po):
    """tests/golden/temporal_48x27_seed11.npz (made by make_golden.py from this oracle) is what the -m gpu test holds the CUDA path to."""
    import os
    import temporal_inputs
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "temporal_48x27_seed11.npz"))
    state = {"acc": np.zeros((27, 48, 4), np.float32), "vel": np.zeros((27, 48, 2), np.float32)}

    def update(data, pos, ids, inv, last):
        state["acc"], vel, ran = po.state_update(data, pos, ids, inv, last, state["acc"])
        if ran:
            state["vel"] = vel

    seq = temporal_inputs.run_sequence(update, lambda: state["acc"].copy(), po, 48, 27, 5, seed=11)
    from conftest import same_bits_or_nan
    assert same_bits_or_nan(np.stack(seq), g["acc_bits"].view(np.float32))
    assert same_bits_or_nan(state["vel"], g["vel_bits"].view(np.float32))
    # the sequence exercises both branches of compute.wgsl and produces NaNs from the poisoned texels without spreading them everywhere
    last = seq[-1]
    assert 0 < np.isnan(last).any(-1).mean() < 0.5
