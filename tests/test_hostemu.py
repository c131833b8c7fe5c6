"""CPU unit tests of the product's __host__ __device__ code (tests/hostemu): traversal and wavefront bodies compiled for
the host and compared with the oracle.  This is a logic check that runs in the GPU-less container; the parity tests
proper are the `gpu` ones that call the CUDA kernels through the C ABI."""
import numpy as np
import pytest

from conftest import bits, compare_closest, same_bits_or_nan


@pytest.fixture(scope="module")
def emu():
    import emu as _emu
    _emu.lib()
    return _emu


def _batches(po, scenes, sd, n):
    st = scenes.pixel_centre_st(sd.width, sd.height, stride=max(1, int((sd.width * sd.height / n) ** 0.5)))
    o, d = po.camera_rays(sd.cam_origin, sd.cam_target, sd.fov_deg, sd.aspect, st)
    yield "primary", o, d, np.full(o.shape[0], np.inf, np.float32)
    yield ("incoherent",) + scenes.incoherent_rays(sd, n)
    yield ("incoherent_finite",) + scenes.incoherent_rays(sd, n, seed=99, finite_tmax=True)


@pytest.mark.parametrize("name", ["c2", "c1", "c3_small", "materials", "c4_small"])
def test_traversal_matches_oracle(po, scenes, emu, name):
    sd = {"c1": scenes.scene_c1, "c2": scenes.scene_c2, "c3_small": lambda: scenes.scene_c3(grid=8), "materials": scenes.scene_materials_test,
          "c4_small": lambda: scenes.scene_c4(nverts=150, n_spheres=8)}[name]()
    sc = po.OracleScene(sd)
    es = emu.EmuScene(sd, po.pack_scene)
    ties = 0
    for _, o, d, tm in _batches(po, scenes, sd, 30000):
        for which in (0, 1):
            ref = sc.trace_closest(o, d, tm, which=which)
            ref_any = sc.trace_any(o, d, tm, which=which)
            # reference-order kernel: bit-exact, no ties allowed
            got = es.trace(o, d, tm, which=which, kernel=1)
            compare_closest(ref, got, sc, o, d, tm, which=which, allow_ties=False)
            assert np.array_equal(es.trace(o, d, tm, which=which, kernel=1, any_hit=True)["hit"], ref_any["hit"])
            # 8-wide kernel: bit-exact up to classified ties
            got = es.trace(o, d, tm, which=which, kernel=0)
            ties += compare_closest(ref, got, sc, o, d, tm, which=which)[1]
            assert np.array_equal(es.trace(o, d, tm, which=which, kernel=0, any_hit=True)["hit"], ref_any["hit"])
    assert ties < 200


@pytest.mark.parametrize("name", ["c2", "c1", "c3_small", "materials", "c4_small"])
def test_multiray_stage_logic_matches_one_ray_traversal(po, scenes, emu, name):
    """multiray.cuh (K rays per lane, gated stages, ray records addressed through MrRay) on one emulated warp against the one-ray
    traversal of traverse.h: the per-ray arithmetic and visit order are the same, so every output is bit-identical, whatever K and the
    quorums — and the oracle's any-hit answers are reproduced."""
    sd = {"c1": scenes.scene_c1, "c2": scenes.scene_c2, "c3_small": lambda: scenes.scene_c3(grid=8), "materials": scenes.scene_materials_test,
          "c4_small": lambda: scenes.scene_c4(nverts=150, n_spheres=8)}[name]()
    sc = po.OracleScene(sd)
    es = emu.EmuScene(sd, po.pack_scene)
    for _, o, d, tm in _batches(po, scenes, sd, 6000):
        for which in (0, 1):
            ref = es.trace(o, d, tm, which=which, kernel=0)
            ref_any = sc.trace_any(o, d, tm, which=which)
            for K, fq, qn, ql, qi in ((1, 24, 8, 1, 1), (2, 24, 12, 8, 8), (3, 8, 1, 1, 1), (4, 100, 32, 32, 32)):
                got = emu.trace_mr(es, o, d, tm, which=which, K=K, free_quorum=fq, qn=qn, ql=ql, qi=qi)
                assert np.array_equal(got["hit"], ref["hit"]), (name, which, K)
                m = ref["hit"] > 0
                for key in ("prim", "inst"):
                    assert np.array_equal(got[key][m], ref[key][m]), (name, which, K, key)
                for key in ("t", "u", "v"):
                    assert np.array_equal(bits(got[key][m]), bits(ref[key][m])), (name, which, K, key)
                ga = emu.trace_mr(es, o, d, tm, which=which, any_hit=True, K=K, free_quorum=fq, qn=qn, ql=ql, qi=qi)
                assert np.array_equal(ga["hit"], ref_any["hit"]), (name, which, K, "any")


@pytest.mark.parametrize("name,W,H", [("c3_small", 96, 56), ("c3_small", 32, 16), ("c4_small", 96, 56), ("c2", 64, 36), ("c1", 64, 36), ("materials", 64, 36)])
def test_tile_lists_hold_every_instance_a_camera_ray_hits(po, scenes, emu, name, W, H):
    """tile_list_build (wavefront.h, the body of k_tile_lists): the list of an 8 x 4 pixel tile must name every TLAS leaf a camera ray of
    the tile can hit — rays through every pixel of the tile at random and at extreme sub-pixel jitter are traced the ordinary way, and the
    instance each one hits has to be in its tile's list with a distance bound not beyond the hit; a tile with an empty list has no hit."""
    sd = {"c1": scenes.scene_c1, "c2": scenes.scene_c2, "c3_small": lambda: scenes.scene_c3(grid=8), "materials": scenes.scene_materials_test,
          "c4_small": lambda: scenes.scene_c4(nverts=150, n_spheres=8)}[name]()
    es = emu.EmuScene(sd, po.pack_scene)
    lists, s2p, leaf = emu.tile_lists(es, sd, W, H)
    assert np.array_equal(np.sort(s2p), np.arange(W * H, dtype=np.uint32)), "the tiled slot layout is a bijection onto the pixels"
    npix = W * H
    pixel = s2p.astype(np.int64)
    x, y = (pixel % W).astype(np.float32), (pixel // W).astype(np.float32)
    tile = np.arange(npix) >> 5
    rng = np.random.default_rng(7)
    n_listed = 0
    for jitter in ("random", "random", "lo", "hi", "mixed"):
        if jitter == "random":
            jx, jy = rng.random(npix, dtype=np.float32) - np.float32(0.5), rng.random(npix, dtype=np.float32) - np.float32(0.5)
        else:
            e = np.float32(0.49999997)
            jx = np.full(npix, -0.5 if jitter == "lo" else e, np.float32)
            jy = np.full(npix, -0.5 if jitter in ("lo", "mixed") else e, np.float32)
        st = np.stack([(x + jx) / np.float32(W), (y + jy) / np.float32(H)], axis=1).astype(np.float32)  # main.rs:196-197 as regen_slot computes it
        o, d = emu.camera_rays(sd.cam_origin, sd.cam_target, sd.fov_deg, np.float32(W) / np.float32(H), st)
        r = es.trace(o, d, np.full(npix, np.inf, np.float32))
        # (a) the traversal that starts from the lists finds what the TLAS walk finds, bit for bit
        g = emu.trace_list(es, lists, tile, o, d)
        assert np.array_equal(g["hit"], r["hit"]), f"{name} {W}x{H} jitter {jitter}: hit / miss differs"
        m = r["hit"] > 0
        for key in ("prim", "inst"):
            assert np.array_equal(g[key][m], r[key][m]), (name, W, H, jitter, key)
        for key in ("t", "u", "v"):
            assert np.array_equal(bits(g[key][m]), bits(r[key][m])), (name, W, H, jitter, key)
        # (b) the lists themselves: the instance a ray hits is named, with a distance bound not beyond the hit
        for j in np.nonzero(r["hit"])[0]:
            row = lists[tile[j]]
            n = int(row[0, 0])
            if n == 0xffffffff:
                continue  # more than a list holds: such a tile walks the TLAS
            entries = row[1:1 + n]
            is_inst = np.ones(n, bool)  # instance entries are followed by their k node entries
            i = 0
            while i < n:
                k = int(entries[i, 0] >> 28)
                is_inst[i + 1:i + 1 + k] = False
                i += 1 + k
            want = leaf[r["inst"][j]]
            k = np.nonzero(is_inst & ((entries[:, 0] & 0x0fffffff) == want))[0]
            assert k.size == 1, f"{name} {W}x{H}: tile {tile[j]} lacks the instance its ray (slot {j}, jitter {jitter}) hits"
            assert entries[k[0], 1].view(np.float32) <= r["t"][j], "the entry-distance bound of a listed leaf lies beyond a hit inside it"
            n_listed += 1
    assert n_listed > 0 or name == "c2"


def test_integer_tier_bit_exact(po, emu):
    idx = (np.arange(20000, dtype=np.uint32) * 7) % 5
    seeds = (np.arange(20000, dtype=np.uint32) * np.uint32(2654435761)).astype(np.uint32)
    assert np.array_equal(bits(po.ss_sobol(idx, seeds)), bits(emu.ss_sobol(idx, seeds)))
    for s, p, k in ((1, 0, 0), (77, 12345, 9), (2**63 + 5, 2**31 + 1, 4095)):
        assert np.array_equal(po.rng_stream(s, p, k, 256), emu.rng_stream(s, p, k, 256))


def test_camera_rays(po, scenes, emu):
    sd = scenes.scene_c2()
    st = scenes.pixel_centre_st(320, 180)
    o1, d1 = po.camera_rays(sd.cam_origin, sd.cam_target, sd.fov_deg, np.float32(320 / 180), st)
    o2, d2 = emu.camera_rays(sd.cam_origin, sd.cam_target, sd.fov_deg, np.float32(320 / 180), st)
    assert np.array_equal(bits(o1), bits(o2)) and np.array_equal(bits(d1), bits(d2))  # camera.rs is in the bit-exact tier


@pytest.mark.parametrize("name,nee", [("c2", True), ("c2", False), ("c1", True), ("materials", True)])
def test_wavefront_matches_integrate_path_by_path(po, scenes, emu, name, nee):
    """Same RNG streams -> the wavefront state machine must trace the same paths as the recursive-loop restatement."""
    sd = {"c1": scenes.scene_c1, "c2": scenes.scene_c2, "materials": scenes.scene_materials_test}[name]()
    sc = po.OracleScene(sd)
    es = emu.EmuScene(sd, po.pack_scene)
    W, H, N = 64, 36, 6
    a = sc.render(width=W, height=H, sample_count=N, seed=5, enable_nee=nee)
    b = emu.render(es, sd, W, H, sample_count=N, seed=5, enable_nee=nee)
    for k in ("extend", "shadow", "probe", "paths"):
        assert abs(a["counters"][k] - b["counters"][k]) <= 0.002 * a["counters"][k] + 2
    assert np.array_equal(a["id"], b["id"])
    A, B = a["accum"][..., :3] / N, b["accum"][..., :3] / N
    rel = np.abs(A - B).max(axis=2) / (np.abs(A).max(axis=2) + 1e-3)
    assert (rel > 1e-3).mean() < 0.01
    assert np.all(b["accum"][..., 3] == N)
    # splitting the sample range reproduces the same sums (streams are keyed by global sample index)
    c1 = emu.render(es, sd, W, H, sample_begin=0, sample_count=2, seed=5, enable_nee=nee)
    c2 = emu.render(es, sd, W, H, sample_begin=2, sample_count=4, seed=5, enable_nee=nee)
    assert np.allclose(c1["accum"] + c2["accum"], b["accum"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("W,H", [(1, 1), (2, 3), (33, 9), (67, 45)])
def test_temporal_bodies_match_oracle(po, emu, W, H):
    """temporal.cuh (the per-texel bodies of k_resolve) on the host against the oracle's reading of the WGSL passes: bit-exact over a
    five-frame sequence with NaN / inf / behind-the-camera texels."""
    import temporal_inputs
    es = {"acc": np.zeros((H, W, 4), np.float32), "vel": np.zeros((H, W, 2), np.float32)}
    os_ = {"acc": np.zeros((H, W, 4), np.float32), "vel": np.zeros((H, W, 2), np.float32)}

    def emu_update(data, pos, ids, inv, last):
        es["acc"], vel = emu.state_update(data, pos, ids, inv, last, es["acc"])
        if vel is not None:
            es["vel"] = vel

    def po_update(data, pos, ids, inv, last):
        os_["acc"], vel, ran = po.state_update(data, pos, ids, inv, last, os_["acc"])
        if ran:
            os_["vel"] = vel

    got = temporal_inputs.run_sequence(emu_update, lambda: (es["acc"].copy(), es["vel"].copy()), po, W, H, 5, seed=W * 7 + H)
    want = temporal_inputs.run_sequence(po_update, lambda: (os_["acc"].copy(), os_["vel"].copy()), po, W, H, 5, seed=W * 7 + H)
    for f, ((ga, gv), (wa, wv)) in enumerate(zip(got, want)):
        assert same_bits_or_nan(ga, wa), f"frame {f}"
        assert same_bits_or_nan(gv, wv), f"frame {f}"


def test_zero_component_rays_through_identity_instances(po, scenes, emu):
    """Identity instances hand the world ray to the BLAS unchanged only when no component of origin or direction is a zero; a zero
    takes the full arithmetic of Ray::transform (ray.rs:22-28: x*1 + y*0 + z*0 + 0 decides its sign, and with it 1/d = +-inf).
    Axis-parallel rays with zeros of either sign, origins with zero coordinates: bit-exact against the oracle, no ties allowed for
    the hit flag and t."""
    sd = scenes.scene_c2()  # every instance is the identity
    sc = po.OracleScene(sd)
    es = emu.EmuScene(sd, po.pack_scene)
    rng = np.random.default_rng(7)
    n = 6000
    o = rng.uniform(-250, 250, size=(n, 3)).astype(np.float32)
    o[:, 1] += 50.0
    d = rng.normal(size=(n, 3)).astype(np.float32)
    zero = rng.integers(0, 8, size=n)  # which direction components are zeroed (never all three)
    for k in range(3):
        m = ((zero >> k) & 1).astype(bool) & (zero != 7)
        d[m, k] = np.where(rng.integers(0, 2, size=m.sum()) == 1, np.float32(-0.0), np.float32(0.0))
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    oz = rng.integers(0, 4, size=n)
    for k in range(3):
        m = oz == k
        o[m, k] = np.where(rng.integers(0, 2, size=m.sum()) == 1, np.float32(-0.0), np.float32(0.0))
    tm = np.full(n, np.inf, np.float32)
    ref = sc.trace_closest(o, d, tm)
    for kernel in (1, 0):
        got = es.trace(o, d, tm, kernel=kernel)
        compare_closest(ref, got, sc, o, d, tm, allow_ties=(kernel == 0))
        assert np.array_equal(es.trace(o, d, tm, kernel=kernel, any_hit=True)["hit"], sc.trace_any(o, d, tm)["hit"])
    assert ref["hit"].mean() > 0.5
