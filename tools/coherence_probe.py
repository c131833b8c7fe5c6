"""How much would ray reordering buy?  Second-bounce rays of a scene (cosine-ish directions about the primary hits' normals) traced as
a resident batch in pixel order, in random order and sorted by (direction octant, Morton code of the origin).  Tuning aid."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import path_tracer_b200 as ptb
from path_tracer_b200 import scenes

def part1by2(x):
    x = x.astype(np.uint64) & 0x3ff
    x = (x | (x << 16)) & 0x30000ff; x = (x | (x << 8)) & 0x300f00f; x = (x | (x << 4)) & 0x30c30c3; x = (x | (x << 2)) & 0x9249249
    return x

for name in sys.argv[1:] or ["c3", "c2"]:
    sd = {"c2": scenes.scene_c2, "c3": scenes.scene_c3, "c4": scenes.scene_c4}[name]()
    ctx = ptb.Context(0); ctx.upload(sd)
    W, H = 1920, 1080
    ctx.set_camera(sd.cam_origin, sd.cam_target, sd.fov_deg, np.float32(W) / np.float32(H))
    st = scenes.pixel_centre_st(W, H)
    o, d = ctx.camera_rays(st)
    tm = np.full(o.shape[0], np.inf, np.float32)
    r = ctx.trace_closest(o, d, tm)
    ctx.trace_upload(o, d, tm)
    ms = min(ctx.trace_resident(repeats=3) for _ in range(2)) / 3
    print(f"{name} primary (pixel order): {o.shape[0] / ms / 1e3:.0f} Mrays/s  hit {r['hit'].mean():.3f}")
    m = r["hit"] > 0
    P = o[m] + d[m] * r["t"][m, None]; N = r["normal"][m]
    rng = np.random.default_rng(1)
    bounce = 0
    while True:
        v = rng.normal(size=P.shape).astype(np.float32); v /= np.linalg.norm(v, axis=1, keepdims=True)
        D = N + v * 0.999; D = (D / np.linalg.norm(D, axis=1, keepdims=True)).astype(np.float32)
        O = (P + N * 1e-2).astype(np.float32)
        n = O.shape[0]; tmx = np.full(n, np.inf, np.float32)
        bounce += 1
        lo, hi = O.min(0), O.max(0)
        q = np.clip(((O - lo) / (hi - lo + 1e-6) * 1023).astype(np.int64), 0, 1023)
        mort = part1by2(q[:, 0]) | (part1by2(q[:, 1]) << 1) | (part1by2(q[:, 2]) << 2)
        octant = ((D[:, 0] < 0).astype(np.uint64) << 2) | ((D[:, 1] < 0).astype(np.uint64) << 1) | (D[:, 2] < 0).astype(np.uint64)
        orders = {"pixel order": np.arange(n), "random": rng.permutation(n), "morton(origin)": np.argsort(mort, kind="stable"),
                  "octant|morton": np.argsort((octant << 30) | mort, kind="stable"), "morton|octant": np.argsort((mort << 3) | octant, kind="stable")}
        for label, idx in orders.items():
            ctx.trace_upload(O[idx], D[idx], tmx)
            ms = min(ctx.trace_resident(repeats=3) for _ in range(2)) / 3
            print(f"{name} bounce {bounce} {label:16s}: {n / ms / 1e3:.0f} Mrays/s ({n} rays)")
        if bounce == 2: break
        rr = ctx.trace_closest(O, D, tmx)
        mm = rr["hit"] > 0
        P = O[mm] + D[mm] * rr["t"][mm, None]; N = rr["normal"][mm]
