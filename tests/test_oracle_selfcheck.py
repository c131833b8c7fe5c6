"""Self-checks of the CPU oracle (nothing upstream pins it, so it must pin itself)."""
import numpy as np

from conftest import bits


def test_sobol_known_answers(po):
    # sampling.rs:33-48: x = bit reversal, y = XOR of DIRECTIONS over the set bits (sampling.rs:4-8, :24-30)
    t = po.sobol_table()
    i = np.arange(512, dtype=np.uint32)
    rev = np.array([int(f"{int(x):032b}"[::-1], 2) for x in i], dtype=np.uint32)
    assert np.array_equal(t[:, 0], rev)
    assert t[:4, 1].tolist() == [0, 0x80000000, 0xC0000000, 0x40000000]
    directions = [0x80000000, 0xc0000000, 0xa0000000, 0xf0000000, 0x88000000, 0xcc000000, 0xaa000000, 0xff000000, 0x80800000]
    for b, dv in enumerate(directions):
        assert t[1 << b, 1] == dv


def test_hashes_known_answers(po):
    L = po.lib()
    # independent python evaluation of sampling.rs:53-91
    def lk(x, seed):
        M = 0xFFFFFFFF
        x ^= (x * 0x3d20adea) & M
        x = (x + seed) & M
        x = (x * ((seed >> 16) | 1)) & M
        x ^= (x * 0x05526c56) & M
        x ^= (x * 0x53a22864) & M
        return x

    def lb(x):
        M = 0xFFFFFFFF
        x ^= x >> 16
        x = (x * 0x21f0aaad) & M
        x ^= x >> 15
        x = (x * 0xd35a2d97) & M
        x ^= x >> 15
        return x

    rng = np.random.default_rng(1)
    for x, s in rng.integers(0, 2**32, size=(200, 2), dtype=np.uint64):
        x, s = int(x), int(s)
        assert L.po_lk_hash(x, s) == lk(x, s)
        assert L.po_low_bias_hash(x) == lb(x)
    # scramble_base2 is a bijection on u32 (checked on a window) and the jitter stays inside [0,1]
    xs = np.arange(1 << 16, dtype=np.uint32)
    ys = np.array([L.po_scramble_base2(int(x), 12345) for x in xs[:4096]], dtype=np.uint32)
    assert len(np.unique(ys)) == len(ys)
    p = po.ss_sobol(xs % 3, (xs * np.uint32(2654435761)).astype(np.uint32))
    assert p.min() >= 0.0 and p.max() <= 1.0
    assert abs(p.mean() - 0.5) < 0.01


def test_traversal_matches_brute_force(po, scenes):
    for sd in (scenes.scene_c2(), scenes.scene_c3(grid=4), scenes.scene_materials_test()):
        sc = po.OracleScene(sd)
        o, d, tm = scenes.incoherent_rays(sd, 20000)
        for which in (0, 1):
            r = sc.trace_closest(o, d, tm, which=which)
            b = sc.trace_brute(o, d, tm, which=which)
            assert np.array_equal(r["hit"], b["hit"])
            h = r["hit"] == 1
            mism = h & ((r["prim"] != b["prim"]) | (r["inst"] != b["inst"]))
            # different visit order may only flip near-ties
            ia, ib = bits(r["t"]).astype(np.int64), bits(b["t"]).astype(np.int64)
            assert np.all(np.abs(ia - ib)[h] <= 1)
            assert mism.sum() <= 0.002 * max(1, h.sum())
            a = sc.trace_any(o, d, tm, which=which)
            assert np.array_equal(a["hit"], r["hit"])  # t_max = inf: any-hit == closest-hit existence


def test_bvh_invariants(po, scenes):
    sd = scenes.scene_c3(grid=4)
    sc = po.OracleScene(sd)
    for b in range(len(sd.blas)):
        d = sc.blas_dump(b)
        meta, boxes = d["meta"], d["boxes"]
        n = meta.shape[0]
        assert d["root"] == n - 1  # post-order: root allocated last (blas_bvh.rs:130-133)
        seen = np.zeros(sd.blas[b][1], np.int32)
        for i in range(n):
            k, a, c = meta[i]
            if k == 0:
                assert a < i and c < i and a < c  # left subtree, right subtree, then parent
                assert np.array_equal(boxes[i, :3], np.minimum(boxes[a, :3], boxes[c, :3]))
                assert np.array_equal(boxes[i, 3:], np.maximum(boxes[a, 3:], boxes[c, 3:]))
            elif k == 1:
                seen[a] += 1
            else:
                for p in d["leaf_prims"][a:a + c]:
                    seen[p] += 1
        assert np.all(seen == 1)
    t = sc.tlas_dump(0)
    nleaf = len(sd.instances)
    assert np.all(t["meta"][:nleaf, 0] == 1) and np.all(t["meta"][nleaf:, 0] == 0)
    assert t["meta"].shape[0] == 2 * nleaf - 1 and t["root"] == 2 * nleaf - 2


def test_white_furnace_and_nee_consistency(po, scenes):
    sd = scenes.scene_furnace(albedo=0.5, le=1.0)
    sc = po.OracleScene(sd)
    n = 64
    im = sc.render(sample_count=n, seed=11)["accum"]
    rgb = im[..., :3] / n
    pos = sc.render(sample_count=1, seed=11)
    m = (pos["id"] & 0xFF) == 1

    def erode(x, k=2):  # keep pixels whose (2k+1)^2 neighbourhood is uniformly inside the mask (jitter crosses silhouettes)
        out = x.copy()
        for dy in range(-k, k + 1):
            for dx in range(-k, k + 1):
                out &= np.roll(np.roll(x, dy, 0), dx, 1)
        return out

    hit_sphere, shell = erode(m), erode(~m)
    # emissive shell seen directly: exactly Le
    assert np.allclose(rgb[shell], 1.0, atol=1e-6)
    # diffuse sphere in a uniform furnace: albedo * Le, MIS(NEE + BSDF) must integrate to it
    assert abs(rgb[hit_sphere].mean() - 0.5) < 0.01
    off = sc.render(sample_count=n, seed=12, enable_nee=False)["accum"][..., :3] / n
    assert abs(off[hit_sphere].mean() - 0.5) < 0.01


def test_nee_on_off_agree_cornell(po, scenes):
    sd = scenes.scene_c2()
    sc = po.OracleScene(sd)
    W, H, n = 96, 54, 192
    a = sc.render(width=W, height=H, sample_count=n, seed=1)["accum"][..., :3].mean() / n
    b = sc.render(width=W, height=H, sample_count=n, seed=2, enable_nee=False)["accum"][..., :3].mean() / n
    assert abs(a - b) / a < 0.05
