"""GPU parity of the temporal pass (State::update, state.rs:505-579: accumulate.wgsl / velocity.wgsl + compute.wgsl as CUDA kernels,
path_tracer_b200/csrc/temporal.cu) against the CPU oracle: bit-exact, NaNs included, through the C ABI."""
import os

import numpy as np
import pytest

import temporal_inputs
from conftest import bits, same_bits_or_nan

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def ptb():
    import path_tracer_b200 as p
    return p


class OracleState:
    def __init__(self, po, W, H):
        self.po, self.acc, self.vel = po, np.zeros((H, W, 4), np.float32), np.zeros((H, W, 2), np.float32)

    def update(self, data, pos, ids, inv, last):
        self.acc, vel, ran = self.po.state_update(data, pos, ids, inv, last, self.acc)
        if ran:
            self.vel = vel


def test_temporal_golden(ptb, po):
    g = np.load(os.path.join(GOLDEN, "temporal_48x27_seed11.npz"))
    ctx = ptb.Context(0)
    ctx.state_reset(48, 27)
    seq = temporal_inputs.run_sequence(ctx.state_update, lambda: ctx.state_read(), po, 48, 27, 5, seed=11)
    assert same_bits_or_nan(np.stack(seq), g["acc_bits"].view(np.float32))
    assert same_bits_or_nan(ctx.state_read(velocity=True)[1], g["vel_bits"].view(np.float32))


@pytest.mark.parametrize("W,H", [(1, 1), (2, 3), (33, 9), (67, 45), (640, 360)])
def test_temporal_sequence_bit_exact(ptb, po, W, H):
    """Five frames (static, moved, moved, static, moved) with NaN / inf / behind-the-camera texels; ragged sizes exercise the clamped
    3x3 window, clamp-to-edge sampling and partially filled 32x8 blocks."""
    ctx = ptb.Context(0)
    ctx.state_reset(W, H)
    st = OracleState(po, W, H)
    got = temporal_inputs.run_sequence(ctx.state_update, lambda: ctx.state_read(velocity=True), po, W, H, 5, seed=W * 7 + H)
    want = temporal_inputs.run_sequence(st.update, lambda: (st.acc.copy(), st.vel.copy()), po, W, H, 5, seed=W * 7 + H)
    for f, ((ga, gv), (wa, wv)) in enumerate(zip(got, want)):
        assert same_bits_or_nan(ga, wa), f"accumulation differs in frame {f}: {np.count_nonzero(bits(ga) != bits(wa))} words"
        assert same_bits_or_nan(gv, wv), f"velocity differs in frame {f}"


def test_temporal_full_size_properties(ptb, po):
    """1920x1080 (BASELINE size): static frames are an exact running sum; a moving frame with no valid history equals the tent-filtered
    input; output alpha is 1 wherever reprojection validated."""
    W, H = 1920, 1080
    rng = np.random.default_rng(5)
    data = rng.random((H, W, 4), np.float32)
    data[..., 3] = 1.0
    pos = np.zeros((H, W, 4), np.float32)
    ids = np.full((H, W), 9, np.uint32)
    m = np.eye(4, dtype=np.float32).reshape(16)
    ctx = ptb.Context(0)
    ctx.state_reset(W, H)
    for _ in range(3):
        ctx.state_update(data, pos, ids, m, m)
    acc = ctx.state_read()
    assert np.array_equal(acc[..., 3], np.full((H, W), 3.0, np.float32))
    assert np.array_equal(acc[..., :3], (data[..., :3] + data[..., :3]) + data[..., :3])
    m2 = m.copy()
    m2[0] = 1.25
    ctx.state_update(data, pos, ids, m2, m)  # previous id is 0 everywhere, current is 9: fallback branch
    acc = ctx.state_read()
    pad = np.pad(data.astype(np.float64), ((1, 1), (1, 1), (0, 0)), mode="edge")
    k = np.array([1.0, 2.0, 1.0]) / 4.0
    want = sum(k[i] * k[j] * pad[i:i + H, j:j + W] for i in range(3) for j in range(3))
    assert np.abs(acc - want).max() < 1e-4


def test_temporal_resident_moving_camera(ptb, po, scenes):
    """The frame loop of main.rs:179-219 with a moving camera, device-resident: ptb_frame_next + ptb_render(1 sample) +
    ptb_state_update_resident per frame, against the oracle's State::update fed with the same frame buffers read back."""
    sd = scenes.scene_c2()
    W, H = 160, 90
    ctx = ptb.Context(0)
    ctx.upload(sd)
    cams = [(sd.cam_origin, sd.cam_target)] * 2 + [((sd.cam_origin[0] + 20.0 * k, sd.cam_origin[1] + 5.0 * k, sd.cam_origin[2]), sd.cam_target) for k in (1, 2, 3)] + \
           [((sd.cam_origin[0] + 60.0, sd.cam_origin[1] + 15.0, sd.cam_origin[2]), sd.cam_target)] * 2
    ctx.set_camera(*cams[0], sd.fov_deg, sd.aspect)
    ctx.frame_begin(W, H)
    ctx.state_reset(W, H)
    st = OracleState(po, W, H)
    last = ctx.view_projection()
    validated = 0
    for f, cam in enumerate(cams):
        ctx.set_camera(*cam, sd.fov_deg, sd.aspect)
        if f:
            ctx.frame_next()
        ctx.render(f, 1, seed=3, max_bounces=sd.max_bounces)
        vp = ctx.view_projection()
        ctx.state_update_resident(vp, last)
        data = ctx.read_accum()
        pos, ids = ctx.read_aux()
        assert np.array_equal(data[..., 3], np.ones((H, W), np.float32))
        if f:
            assert np.array_equal(ids >> 16, prev_ids & 0xFFFF), "id history (main.rs:206) lost between frames"
        prev_ids = ids
        st.update(data, pos, ids, vp, last)
        got = ctx.state_read()
        assert same_bits_or_nan(got, st.acc), f"frame {f}"
        if not np.array_equal(vp, last):
            # share of texels whose reprojection lands inside the image on the same model id (compute.wgsl:156-161), from the velocity texture
            vel = ctx.state_read(velocity=True)[1].astype(np.float64)
            ys, xs = np.mgrid[0:H, 0:W]
            pc = np.floor((np.stack([(xs + 0.5) / W, (ys + 0.5) / H], -1) - vel) * [W, H]).astype(int)
            inside = (pc[..., 0] >= 0) & (pc[..., 0] < W) & (pc[..., 1] >= 0) & (pc[..., 1] < H)
            old = ids[np.clip(pc[..., 1], 0, H - 1), np.clip(pc[..., 0], 0, W - 1)] >> 16
            validated += int(np.count_nonzero(inside & (old == (ids & 0xFFFF))))
        last = vp
    assert validated > 0.8 * W * H * 3  # the moving frames mostly reproject onto the same model: the history branch is what ran
    rgb, ref = ctx.state_read_rgb8().astype(np.int32), po.tonemap_rgb8(st.acc).astype(np.int32)
    assert np.abs(rgb - ref).max() <= 1 and (rgb != ref).mean() < 0.01  # libm vs CUDA powf: 1 LSB (see test_output_stage_rgb8)
