"""ncu target: one launch each of the one-ray-per-lane kernel (0) and the multiray kernels (3: K = 2, 4: K = 3) on the same resident first-bounce rays.
python tools/mr_ncu_target.py c3"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import path_tracer_b200 as ptb
from path_tracer_b200 import scenes

sd = {"c2": scenes.scene_c2, "c3": scenes.scene_c3, "c4": scenes.scene_c4}[sys.argv[1]]()
ctx = ptb.Context(0)
ctx.upload(sd)
ctx.set_camera(sd.cam_origin, sd.cam_target, sd.fov_deg, sd.aspect)
st = scenes.pixel_centre_st(sd.width, sd.height)
po, pd = ctx.camera_rays(st)
inf = np.full(po.shape[0], np.inf, np.float32)
r = ctx.trace_closest(po, pd, inf, kernel=1)
m = r["hit"] > 0
P = po[m] + pd[m] * r["t"][m, None]
N = r["normal"][m]
rng = np.random.default_rng(1)
v = rng.normal(size=P.shape).astype(np.float32)
v /= np.linalg.norm(v, axis=1, keepdims=True)
D = N + v * np.float32(0.999)
D = (D / np.linalg.norm(D, axis=1, keepdims=True)).astype(np.float32)
P = (P + N * np.float32(1e-2)).astype(np.float32)
ctx.trace_upload(P, D, np.full(P.shape[0], np.inf, np.float32))
for k in (0, 3, 4):
    print(k, P.shape[0], ctx.trace_resident(kernel=k, any_hit=False, repeats=1))
