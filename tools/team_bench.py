"""Team traversal (kernel 2: eight lanes per ray) against the one-ray-per-lane persistent kernel (kernel 0) on resident ray batches:
throughput on large batches, latency on small ones.  python tools/team_bench.py c3"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import path_tracer_b200 as ptb
from path_tracer_b200 import scenes

name = sys.argv[1]
sd = {"c2": scenes.scene_c2, "c3": scenes.scene_c3, "c4": scenes.scene_c4}[name]()
ctx = ptb.Context(0)
ctx.upload(sd)
ctx.set_camera(sd.cam_origin, sd.cam_target, sd.fov_deg, sd.aspect)
for label, n in (("incoherent", 1 << 21), ("incoherent", 1 << 16), ("incoherent", 1 << 13), ("incoherent", 1 << 10), ("primary", 1 << 21)):
    if label == "primary":
        st = scenes.pixel_centre_st(sd.width, sd.height, stride=max(1, int((sd.width * sd.height / n) ** 0.5)))[:n]
        o, d = ctx.camera_rays(st)
        tm = np.full(o.shape[0], np.inf, np.float32)
    else:
        o, d, tm = scenes.incoherent_rays(sd, n)
    ctx.trace_upload(o, d, tm)
    rec = dict(scene=name, batch=label, rays=int(o.shape[0]))
    for k, kn in ((0, "lane"), (2, "team")):
        for a in (False, True):
            ctx.trace_resident(kernel=k, any_hit=a, repeats=2)
            ms = ctx.trace_resident(kernel=k, any_hit=a, repeats=10)
            rec[f"{kn}_{'any' if a else 'closest'}_us"] = round(ms * 1e3, 1)
            rec[f"{kn}_{'any' if a else 'closest'}_mrays"] = round(o.shape[0] / ms / 1e3, 1)
    print(json.dumps(rec))
