"""Where a C3-style frame goes, dense launches against thin ones (queue < 4 entries per resident lane), per option set:
python tools/thin_probe.py c3 64 node_quorum=8 node_quorum=-8 "node_quorum=-8,refill_below=4"     (one render each, profiling on)"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import path_tracer_b200 as ptb
from path_tracer_b200 import scenes

name, spp = sys.argv[1], int(sys.argv[2])
sd = {"c2": scenes.scene_c2, "c3": scenes.scene_c3, "c4": scenes.scene_c4}[name]()
ctx = ptb.Context(0)
ctx.upload(sd)
ctx.set_camera(sd.cam_origin, sd.cam_target, sd.fov_deg, sd.aspect)
ctx.frame_begin(sd.width, sd.height)
ctx.render(0, 2, seed=1, max_bounces=sd.max_bounces)
base = dict(node_quorum=8, refill_below=8, nee_refill_below=8)
for spec in sys.argv[3:]:
    opts = dict(base)
    opts.update({k: int(v) for k, v in (kv.split("=") for kv in spec.split(","))})
    for k, v in opts.items():
        ctx.set_option(k, v)
    best = None
    for rep in range(2):
        ctx.frame_begin(sd.width, sd.height)
        ctx.reset_counters()
        ctx.set_profiling(True)
        ctx.render(0, spp, seed=1, max_bounces=sd.max_bounces)
        t = ctx.timing()
        if best is None or t["render"][0] < best[0]["render"][0]:
            best = (t, ctx.launch_log(), ctx.iteration_log(), ctx.counters())
    ctx.set_profiling(False)
    t, (st, us), act, c = best
    n_slots = sd.width * sd.height * min(16, spp)
    rec = dict(opts=spec, render_ms=round(t["render"][0], 2), tail_ms=round(t["tail"][0], 2), iterations=int(len(act)))
    for cls, nm in ((2, "extend"), (3, "shade"), (4, "nee")):
        u = us[st == cls]
        rays_in = np.concatenate([[n_slots], act[: max(len(u) - 1, 0)]]).astype(np.float64)[: len(u)]
        dense = rays_in >= 4 * 148 * 7 * 128
        rec[nm] = dict(dense_ms=round(float(u[dense].sum()) / 1e3, 2), thin_ms=round(float(u[~dense].sum()) / 1e3, 2), thin_launches=int((~dense).sum()))
    print(json.dumps(rec))
