"""Procedural scenes and fixed ray batches.

The reference hard-codes its scene in `src/main.rs:80-127` and loads OBJ files that are not shipped
(`models/cornell/*.obj`), so every configuration of BASELINE.json is generated here, integer-seeded and
rounded through f32 once.  The arrays produced are exactly what `ptb_scene_upload` (include/ptb.h) takes:

* ``verts``      float32 [ntri, 3, 6]   per vertex: position xyz, normal xyz  (`Vertex`, model.rs:13-18)
* ``blas``       list of (first_tri, tri_count, material dict)               (`Model` order, tlas.rs:31-33)
* ``instances``  list of (blas_index, affine[12] column-major)               (`Model.matrices`, model.rs:27-33)

Quirks of the reference that the scenes respect: instance matrices are translations and
90-degree-multiple rotations only (AABB::transform maps two corners, boundingbox.rs:51-57), emitters sit at
identity (light sampling ignores instance matrices, scene.rs:41-44), scale is in the hundreds of units
(EPSILON = 5e-4, utility.rs:4).
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

LAMBERTIAN, EMISSIVE, SPECULAR, GGX_METAL, GGX_DIELECTRIC, DIELECTRIC = range(6)

IDENTITY = (1.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0)


def material(kind, colour, roughness=0.0, ior=1.0, volume=None):
    """Material POD mirroring the constructors of material.rs (Lambertian::new :99, Emissive::new :126,
    Specular::new :146, GGX::new_metal :290, GGX::new_dielectric :305, Dielectric::new :475).
    ``volume`` = (absorption rgb, k, c, g) as in Volume::new (volume.rs:136)."""
    m = dict(kind=int(kind), colour=tuple(float(c) for c in colour), roughness=float(roughness), ior=float(ior), volume=None)
    if volume is not None:
        a, k, c, g = volume
        m["volume"] = (tuple(float(x) for x in a), float(k), float(c), float(g))
    return m


@dataclass
class SceneData:
    name: str
    verts: np.ndarray
    blas: list
    instances: list
    cam_origin: tuple
    cam_target: tuple
    fov_deg: float
    width: int
    height: int
    spp: int
    max_bounces: int
    notes: dict = field(default_factory=dict)

    @property
    def aspect(self):
        return np.float32(self.width) / np.float32(self.height)

    @property
    def ntri(self):
        return int(self.verts.shape[0])

    @property
    def instanced_tris(self):
        return int(sum(self.blas[b][1] for b, _ in self.instances))


# --------------------------------------------------------------------------------------------------------
# mesh helpers (all float32)
# --------------------------------------------------------------------------------------------------------
def _tri_soup(positions, normals=None):
    """positions [n,3,3] -> verts [n,3,6]; flat face normals cross(p1-p0, p2-p0) when none are given
    (the OBJ loader's fallback, blas.rs:112-116 — left un-normalised there as well)."""
    p = np.asarray(positions, dtype=np.float32)
    if normals is None:
        n = np.cross(p[:, 1] - p[:, 0], p[:, 2] - p[:, 0]).astype(np.float32)
        ln = np.linalg.norm(n, axis=1, keepdims=True).astype(np.float32)
        n = (n / np.maximum(ln, np.float32(1e-30))).astype(np.float32)
        n = np.repeat(n[:, None, :], 3, axis=1)
    else:
        n = np.asarray(normals, dtype=np.float32)
    out = np.concatenate([p, n], axis=2).astype(np.float32)
    out += np.float32(0.0)  # canonicalise -0.0 -> +0.0
    return np.ascontiguousarray(out)


def quad(a, b, c, d):
    """Two triangles (a,b,c), (a,c,d) — the fan triangulation of blas.rs:97-119."""
    a, b, c, d = (np.asarray(x, dtype=np.float32) for x in (a, b, c, d))
    return np.stack([np.stack([a, b, c]), np.stack([a, c, d])])


def box_mesh(lo, hi, rot_y_deg=0.0, centre=None):
    """12 triangles, outward facing, optionally rotated about Y around its centre (baked into vertices)."""
    lo = np.asarray(lo, dtype=np.float64)
    hi = np.asarray(hi, dtype=np.float64)
    x0, y0, z0 = lo
    x1, y1, z1 = hi
    c = [(x0, y0, z0), (x1, y0, z0), (x1, y1, z0), (x0, y1, z0), (x0, y0, z1), (x1, y0, z1), (x1, y1, z1), (x0, y1, z1)]
    faces = [(0, 3, 2, 1), (4, 5, 6, 7), (0, 4, 7, 3), (1, 2, 6, 5), (3, 7, 6, 2), (0, 1, 5, 4)]
    pts = np.array(c, dtype=np.float64)
    if rot_y_deg:
        ctr = (lo + hi) / 2 if centre is None else np.asarray(centre, dtype=np.float64)
        th = np.deg2rad(rot_y_deg)
        r = np.array([[np.cos(th), 0, np.sin(th)], [0, 1, 0], [-np.sin(th), 0, np.cos(th)]])
        pts = (pts - ctr) @ r.T + ctr
    pts = pts.astype(np.float32)
    tris = [quad(pts[f[0]], pts[f[1]], pts[f[2]], pts[f[3]]) for f in faces]
    return np.concatenate(tris)


def torus_mesh(R=40.0, r=14.0, nu=32, nv=16):
    """nu*nv quads -> 2*nu*nv triangles with smooth vertex normals; axis = Y."""
    u = (np.arange(nu + 1) % nu) * (2 * np.pi / nu)
    v = (np.arange(nv + 1) % nv) * (2 * np.pi / nv)
    uu, vv = np.meshgrid(u, v, indexing="ij")
    cx, sx = np.cos(uu), np.sin(uu)
    cv, sv = np.cos(vv), np.sin(vv)
    pos = np.stack([(R + r * cv) * cx, r * sv, (R + r * cv) * sx], axis=-1)
    nrm = np.stack([cv * cx, sv, cv * sx], axis=-1)
    return _grid_to_tris(pos, nrm, flip=True)


def _grid_to_tris(pos, nrm, flip=False):
    """pos,nrm [a+1,b+1,3] -> (positions [2ab,3,3], normals [2ab,3,3])"""
    p00, p10, p11, p01 = pos[:-1, :-1], pos[1:, :-1], pos[1:, 1:], pos[:-1, 1:]
    n00, n10, n11, n01 = nrm[:-1, :-1], nrm[1:, :-1], nrm[1:, 1:], nrm[:-1, 1:]
    if flip:
        t1p, t1n = np.stack([p00, p11, p10], axis=-2), np.stack([n00, n11, n10], axis=-2)
        t2p, t2n = np.stack([p00, p01, p11], axis=-2), np.stack([n00, n01, n11], axis=-2)
    else:
        t1p, t1n = np.stack([p00, p10, p11], axis=-2), np.stack([n00, n10, n11], axis=-2)
        t2p, t2n = np.stack([p00, p11, p01], axis=-2), np.stack([n00, n11, n01], axis=-2)
    P = np.stack([t1p, t2p], axis=2).reshape(-1, 3, 3)
    N = np.stack([t1n, t2n], axis=2).reshape(-1, 3, 3)
    return P.astype(np.float32), N.astype(np.float32)


def icosphere(radius=30.0, subdiv=3):
    """20 * 4**subdiv triangles, smooth normals."""
    t = (1.0 + 5.0 ** 0.5) / 2.0
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t], [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], dtype=np.float64)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6], [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]])
    tri = v[f]
    for _ in range(subdiv):
        a, b, c = tri[:, 0], tri[:, 1], tri[:, 2]
        ab, bc, ca = a + b, b + c, c + a
        ab /= np.linalg.norm(ab, axis=1, keepdims=True)
        bc /= np.linalg.norm(bc, axis=1, keepdims=True)
        ca /= np.linalg.norm(ca, axis=1, keepdims=True)
        tri = np.concatenate([np.stack([a, ab, ca], 1), np.stack([b, bc, ab], 1), np.stack([c, ca, bc], 1), np.stack([ab, bc, ca], 1)])
    return (tri * radius).astype(np.float32), tri.astype(np.float32)


# integer hashes of sampling.rs:76-91, vectorised (used for terrain and ray batches)
def low_bias_hash(x):
    x = np.asarray(x, dtype=np.uint32).copy()
    x ^= x >> np.uint32(16)
    x *= np.uint32(0x21F0AAAD)
    x ^= x >> np.uint32(15)
    x *= np.uint32(0xD35A2D97)
    x ^= x >> np.uint32(15)
    return x


def _u01(h):
    return (h.astype(np.float64) / 4294967296.0)


def terrain_mesh(nverts=2237, extent=4000.0, amplitude=300.0, octaves=6, seed=0x5EED0004):
    """Heightfield nverts x nverts -> 2*(nverts-1)^2 triangles; fBm of bilinear value noise over an integer hash
    lattice low_bias_hash(ix, iz, octave)."""
    n = nverts
    xs = np.linspace(-extent / 2, extent / 2, n)
    gx, gz = np.meshgrid(xs, xs, indexing="ij")
    h = np.zeros((n, n), dtype=np.float64)
    amp, freq = 1.0, 4.0
    norm = 0.0
    for o in range(octaves):
        fx = (gx / extent + 0.5) * freq
        fz = (gz / extent + 0.5) * freq
        ix, iz = np.floor(fx).astype(np.int64), np.floor(fz).astype(np.int64)
        tx, tz = fx - ix, fz - iz
        tx = tx * tx * (3 - 2 * tx)
        tz = tz * tz * (3 - 2 * tz)

        def lat(i, j):
            key = (i.astype(np.uint32) * np.uint32(0x9E3779B1)) ^ (j.astype(np.uint32) * np.uint32(0x85EBCA77)) ^ np.uint32((seed + o * 0x632BE5AB) & 0xFFFFFFFF)
            return _u01(low_bias_hash(key))

        v = (lat(ix, iz) * (1 - tx) + lat(ix + 1, iz) * tx) * (1 - tz) + (lat(ix, iz + 1) * (1 - tx) + lat(ix + 1, iz + 1) * tx) * tz
        h += amp * v
        norm += amp
        amp *= 0.5
        freq *= 2.0
    h = (h / norm - 0.5) * 2.0 * amplitude
    pos = np.stack([gx, h, gz], axis=-1)
    dx = np.gradient(h, xs, axis=0)
    dz = np.gradient(h, xs, axis=1)
    nrm = np.stack([-dx, np.ones_like(h), -dz], axis=-1)
    nrm /= np.linalg.norm(nrm, axis=-1, keepdims=True)
    return _grid_to_tris(pos, nrm, flip=True)


def translate(t):
    return (1.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 1.0, float(t[0]), float(t[1]), float(t[2]))


def rot_y90(k, t=(0.0, 0.0, 0.0)):
    """Exact rotation by k*90 degrees about +Y (entries 0/+-1) plus translation, glam column-major."""
    c, s = [(1.0, 0.0), (0.0, 1.0), (-1.0, 0.0), (0.0, -1.0)][k % 4]
    # columns: x_axis = (c, 0, -s), y_axis = (0,1,0), z_axis = (s, 0, c)
    return (c, 0.0, -s, 0.0, 1.0, 0.0, s, 0.0, c, float(t[0]), float(t[1]), float(t[2]))


class _Builder:
    def __init__(self):
        self.chunks, self.blas, self.instances, self.n = [], [], [], 0

    def add_blas(self, verts, mat, matrices=(IDENTITY,)):
        verts = np.ascontiguousarray(verts, dtype=np.float32)
        self.blas.append((self.n, int(verts.shape[0]), mat))
        self.chunks.append(verts)
        self.n += int(verts.shape[0])
        b = len(self.blas) - 1
        for m in matrices:
            self.instances.append((b, tuple(np.float32(x).item() for x in m)))
        return b

    def verts(self):
        return np.ascontiguousarray(np.concatenate(self.chunks, axis=0))


# --------------------------------------------------------------------------------------------------------
# the Cornell shell shared by C1/C2 (materials and camera: main.rs:82-92, 101-104, 122-127)
# --------------------------------------------------------------------------------------------------------
X0, X1, Y0, Y1, Z0, Z1 = -278.0, 278.0, -228.0, 328.0, -278.0, 278.0


def _cornell_shell(b: _Builder):
    light = quad((-65, Y1 - 0.5, -65), (65, Y1 - 0.5, -65), (65, Y1 - 0.5, 65), (-65, Y1 - 0.5, 65))
    b.add_blas(_tri_soup(light), material(EMISSIVE, (15, 15, 15)))                                    # cb_light, main.rs:101
    main = np.concatenate([
        quad((X0, Y0, Z0), (X0, Y0, Z1), (X1, Y0, Z1), (X1, Y0, Z0)),   # floor (normal +y)
        quad((X0, Y1, Z0), (X1, Y1, Z0), (X1, Y1, Z1), (X0, Y1, Z1)),   # ceiling (normal -y)
        quad((X0, Y0, Z0), (X1, Y0, Z0), (X1, Y1, Z0), (X0, Y1, Z0)),   # back wall (normal +z)
    ])
    b.add_blas(_tri_soup(main), material(LAMBERTIAN, (0.73, 0.73, 0.73)))                             # cb_main, main.rs:102
    right = quad((X1, Y0, Z0), (X1, Y0, Z1), (X1, Y1, Z1), (X1, Y1, Z0))
    b.add_blas(_tri_soup(right), material(LAMBERTIAN, (0.65, 0.05, 0.05)))                            # cb_right, main.rs:103
    left = quad((X0, Y0, Z0), (X0, Y1, Z0), (X0, Y1, Z1), (X0, Y0, Z1))
    b.add_blas(_tri_soup(left), material(LAMBERTIAN, (0.12, 0.45, 0.15)))                             # cb_left, main.rs:104


def scene_c1(with_dragon=True, width=1024, height=576, spp=64):
    """BASELINE config 0: stand-in for the reference default scene (main.rs:80-127): Cornell shell as 4 BLASes at
    identity plus two instances of a closed mesh with the brown-glass GGX dielectric + homogeneous medium."""
    b = _Builder()
    _cornell_shell(b)
    if with_dragon:
        p, n = torus_mesh(R=70.0, r=32.0, nu=48, nv=24)
        p = p + np.array([0.0, -150.0, 0.0], dtype=np.float32)
        vol = ((0.4, 0.62, 0.7), 0.1, 1.0 / 200.0, 0.6)                                                # main.rs:80
        glass = material(GGX_DIELECTRIC, (0.95, 0.95, 0.95), roughness=0.2, ior=1.5, volume=vol)       # main.rs:87
        # main.rs:97-113: identity and rot-Y(pi) + translate (0,200,0)
        b.add_blas(_tri_soup(p, n), glass, matrices=(IDENTITY, rot_y90(2, (0.0, 200.0, 0.0))))
    return SceneData("c1_default_standin", b.verts(), b.blas, b.instances, (0.0, 50.0, 1000.0), (0.0, 50.0, 0.0), 60.0, width, height, spp, 1024)


def scene_c2(width=1920, height=1080, spp=256, max_bounces=8):
    """BASELINE config 1: synthetic Cornell box, 36 triangles, diffuse + area light."""
    b = _Builder()
    _cornell_shell(b)
    tall = box_mesh((-190.0, Y0, -190.0), (-25.0, Y0 + 330.0, -25.0), rot_y_deg=17.0)
    b.add_blas(_tri_soup(tall), material(LAMBERTIAN, (0.05, 0.05, 0.25)))                             # diffuse_blue, main.rs:85
    short = box_mesh((30.0, Y0, 10.0), (195.0, Y0 + 165.0, 175.0), rot_y_deg=-18.0)
    b.add_blas(_tri_soup(short), material(LAMBERTIAN, (0.73, 0.73, 0.73)))
    return SceneData("c2_cornell36", b.verts(), b.blas, b.instances, (0.0, 50.0, 1000.0), (0.0, 50.0, 0.0), 60.0, width, height, spp, max_bounces)


def scene_c3(width=1920, height=1080, spp=64, grid=32, max_bounces=1024, seed=0x5EED0003):
    """BASELINE config 2: ~1M instanced triangles: two 1024-triangle tori (Lambertian, GGX metal 0.4),
    grid*grid/2 instances each on a jittered grid (pitch 120, rot-Y in {0,90,180,270}), ground and emitter."""
    b = _Builder()
    half = grid * 120.0 / 2 + 200.0
    light = quad((-600, 700, -600), (600, 700, -600), (600, 700, 600), (-600, 700, 600))
    b.add_blas(_tri_soup(light), material(EMISSIVE, (15, 15, 15)))
    ground = quad((-half, 0.0, -half), (-half, 0.0, half), (half, 0.0, half), (half, 0.0, -half))
    b.add_blas(_tri_soup(ground), material(LAMBERTIAN, (0.73, 0.73, 0.73)))
    p, n = torus_mesh(R=40.0, r=14.0, nu=32, nv=16)
    idx = np.arange(grid * grid, dtype=np.uint32)
    hx = _u01(low_bias_hash(idx * np.uint32(4) + np.uint32(seed)))
    hz = _u01(low_bias_hash(idx * np.uint32(4) + np.uint32(seed + 1)))
    hy = _u01(low_bias_hash(idx * np.uint32(4) + np.uint32(seed + 2)))
    hr = low_bias_hash(idx * np.uint32(4) + np.uint32(seed + 3)) >> np.uint32(30)
    mats = {0: [], 1: []}
    for i in range(grid * grid):
        gx, gz = i % grid, i // grid
        t = (np.float32((gx - grid / 2 + 0.5) * 120.0 + (hx[i] - 0.5) * 30.0), np.float32(20.0 + hy[i] * 60.0), np.float32((gz - grid / 2 + 0.5) * 120.0 + (hz[i] - 0.5) * 30.0))
        mats[(gx + gz) & 1].append(rot_y90(int(hr[i]), t))
    b.add_blas(_tri_soup(p, n), material(LAMBERTIAN, (0.6, 0.35, 0.2)), matrices=mats[0])
    b.add_blas(_tri_soup(p, n), material(GGX_METAL, (0.1, 0.1, 0.45), roughness=0.4), matrices=mats[1])  # ggx_blue, main.rs:86
    span = grid * 120.0
    return SceneData("c3_instanced_1m" if grid == 32 else f"c3_instanced_g{grid}", b.verts(), b.blas, b.instances, (0.0, span * 0.22, span * 0.62), (0.0, 20.0, 0.0), 60.0, width, height, spp, max_bounces)


def scene_c4(width=3840, height=2160, spp=16, nverts=2237, max_bounces=16, n_spheres=64, seed=0x5EED0004):
    """BASELINE config 3: displaced terrain (2*(nverts-1)^2 triangles, ~10M at 2237) with dielectric and GGX-metal
    icospheres above it and an emitter quad."""
    b = _Builder()
    light = quad((-900, 1400, -900), (900, 1400, -900), (900, 1400, 900), (-900, 1400, 900))
    b.add_blas(_tri_soup(light), material(EMISSIVE, (40, 38, 34)))
    p, n = terrain_mesh(nverts=nverts, seed=seed)
    b.add_blas(_tri_soup(p, n), material(LAMBERTIAN, (0.45, 0.42, 0.35)))
    sp, sn = icosphere(radius=45.0, subdiv=3)
    idx = np.arange(2 * n_spheres, dtype=np.uint32)
    hx = _u01(low_bias_hash(idx * np.uint32(3) + np.uint32(seed + 11)))
    hz = _u01(low_bias_hash(idx * np.uint32(3) + np.uint32(seed + 12)))
    hy = _u01(low_bias_hash(idx * np.uint32(3) + np.uint32(seed + 13)))
    tr = [translate((np.float32((hx[i] - 0.5) * 3000.0), np.float32(380.0 + hy[i] * 250.0), np.float32((hz[i] - 0.5) * 3000.0))) for i in range(2 * n_spheres)]
    b.add_blas(_tri_soup(sp, sn), material(DIELECTRIC, (0.95, 0.95, 0.95), ior=1.5), matrices=tr[:n_spheres])   # glass, main.rs:89
    b.add_blas(_tri_soup(sp, sn), material(GGX_METAL, (0.8, 0.6, 0.3), roughness=0.3), matrices=tr[n_spheres:])
    name = "c4_terrain_10m" if nverts == 2237 else f"c4_terrain_v{nverts}"
    return SceneData(name, b.verts(), b.blas, b.instances, (0.0, 900.0, 2600.0), (0.0, 150.0, 0.0), 55.0, width, height, spp, max_bounces)


def scene_materials_test(width=256, height=144, spp=16):
    """Small scene exercising every material kind (mirror, smooth glass with medium, rough glass, metal)."""
    b = _Builder()
    _cornell_shell(b)
    sp, sn = icosphere(radius=70.0, subdiv=2)
    vol = ((0.4, 0.62, 0.7), 0.1, 1.0 / 200.0, 0.6)
    b.add_blas(_tri_soup(sp + np.float32([-150, -158, 60]), sn), material(DIELECTRIC, (0.95, 0.95, 0.95), ior=1.5, volume=vol))
    b.add_blas(_tri_soup(sp + np.float32([0, -158, -120]), sn), material(SPECULAR, (1, 1, 1)))
    b.add_blas(_tri_soup(sp, sn), material(GGX_METAL, (0.1, 0.1, 0.45), roughness=0.4), matrices=(translate((150, -158, 60)),))
    b.add_blas(_tri_soup(sp, sn), material(GGX_DIELECTRIC, (1, 1, 1), roughness=0.2, ior=1.5), matrices=(rot_y90(1, (0.0, 60.0, 0.0)),))
    return SceneData("materials_test", b.verts(), b.blas, b.instances, (0.0, 50.0, 1000.0), (0.0, 50.0, 0.0), 60.0, width, height, spp, 64)


def scene_furnace(albedo=0.5, le=1.0, width=64, height=64, spp=64):
    """Closed Lambertian box whose every surface... cannot emit and scatter at once upstream (Emissive terminates),
    so the furnace is a closed emissive box seen through nothing: every path hits Le on its first bounce -> radiance Le;
    plus a Lambertian sphere inside: with albedo a and uniform Le the radiance leaving the sphere is a*Le."""
    b = _Builder()
    shell = box_mesh((-300, -300, -300), (300, 300, 1300))
    shell = shell[:, ::-1, :]  # inward facing (irrelevant to emission)
    b.add_blas(_tri_soup(shell), material(EMISSIVE, (le, le, le)))
    sp, sn = icosphere(radius=120.0, subdiv=3)
    b.add_blas(_tri_soup(sp, sn), material(LAMBERTIAN, (albedo, albedo, albedo)))
    return SceneData("furnace", b.verts(), b.blas, b.instances, (0.0, 0.0, 1000.0), (0.0, 0.0, 0.0), 20.0, width, height, spp, 64)


# --------------------------------------------------------------------------------------------------------
# fixed ray batches: deterministic, hash-generated
# --------------------------------------------------------------------------------------------------------
def scene_bounds(sd: SceneData):
    """World AABB of all instances (conservative: 8 corners of each BLAS box through its matrix)."""
    lo, hi = np.full(3, np.inf), np.full(3, -np.inf)
    for b, m in sd.instances:
        first, cnt, _ = sd.blas[b]
        p = sd.verts[first:first + cnt, :, :3].reshape(-1, 3)
        bl, bh = p.min(0).astype(np.float64), p.max(0).astype(np.float64)
        M = np.array(m[:9], dtype=np.float64).reshape(3, 3).T
        t = np.array(m[9:], dtype=np.float64)
        for cx in (bl[0], bh[0]):
            for cy in (bl[1], bh[1]):
                for cz in (bl[2], bh[2]):
                    q = M @ np.array([cx, cy, cz]) + t
                    lo, hi = np.minimum(lo, q), np.maximum(hi, q)
    return lo, hi


def incoherent_rays(sd: SceneData, n, seed=0x1234, finite_tmax=False):
    """origin uniform in the scene AABB, direction uniform on the sphere, from low_bias_hash(index*8+k)."""
    lo, hi = scene_bounds(sd)
    idx = np.arange(n, dtype=np.uint32) * np.uint32(8) + np.uint32(seed)
    u = [_u01(low_bias_hash(idx + np.uint32(k))) for k in range(6)]
    o = np.stack([lo[k] + (hi[k] - lo[k]) * u[k] for k in range(3)], axis=1)
    z = 1.0 - 2.0 * u[3]
    phi = 2.0 * np.pi * u[4]
    r = np.sqrt(np.maximum(0.0, 1.0 - z * z))
    d = np.stack([r * np.cos(phi), z, r * np.sin(phi)], axis=1)
    d32 = d.astype(np.float32)
    d32 = normalize_f32(d32)
    if finite_tmax:
        diag = float(np.linalg.norm(hi - lo))
        tmax = (u[5] * diag).astype(np.float32)
    else:
        tmax = np.full(n, np.inf, dtype=np.float32)
    return np.ascontiguousarray(o.astype(np.float32)), np.ascontiguousarray(d32), tmax


def normalize_f32(d):
    """Normalise in f32 (directions handed to Ray::new must be unit length, ray.rs:12)."""
    d = d.astype(np.float32)
    l = np.sqrt((d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]).astype(np.float32)
    return np.ascontiguousarray((d / l[:, None]).astype(np.float32))


def pixel_centre_st(width, height, n=None, stride=1):
    """(s,t) of pixel centres as main.rs:196-197 with zero jitter: u = x/W, v = y/H."""
    ys, xs = np.meshgrid(np.arange(0, height, stride, dtype=np.float32), np.arange(0, width, stride, dtype=np.float32), indexing="ij")
    st = np.stack([xs.ravel() / np.float32(width), ys.ravel() / np.float32(height)], axis=1).astype(np.float32)
    if n is not None:
        st = st[:n]
    return np.ascontiguousarray(st)
