"""Option sweep on one scene (tuning aid, not the contract bench): python tools/sweep.py c3 8 extend_blocks=5,6,7,8 refill_below=8,16,24"""
import itertools
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import path_tracer_b200 as ptb
from path_tracer_b200 import api as _api
if os.environ.get("PTB_LIB"):  # A/B runs of two builds in one gpurun call (tuning aid)
    _api.lib_path = lambda: os.environ["PTB_LIB"]
from path_tracer_b200 import scenes

name, spp = sys.argv[1], int(sys.argv[2])
axes = [(a.split("=")[0], [int(v) for v in a.split("=")[1].split(",")]) for a in sys.argv[3:]]
sd = {"c1": scenes.scene_c1, "c2": scenes.scene_c2, "c3": scenes.scene_c3, "c4": scenes.scene_c4, "c4s": lambda: scenes.scene_c4(nverts=700)}[name]()
ctx = ptb.Context(0)
ctx.upload(sd)
W, H = sd.width, sd.height
ctx.set_camera(sd.cam_origin, sd.cam_target, sd.fov_deg, sd.aspect)
ctx.frame_begin(W, H)
ctx.render(0, 2, seed=1, max_bounces=sd.max_bounces)
o, d, tm = scenes.incoherent_rays(sd, 1 << 21)
ctx.trace_upload(o, d, tm)
for combo in itertools.product(*[v for _, v in axes]):
    opts = dict(zip([k for k, _ in axes], combo))
    for k, v in opts.items():
        ctx.set_option(k, v)
    best = None
    for rep in range(2):
        ctx.frame_begin(W, H)
        ctx.reset_counters()
        ctx.set_profiling(not os.environ.get("PTB_NOPROF"))
        ctx.render(0, spp, seed=1, max_bounces=sd.max_bounces)
        t, c = ctx.timing(), ctx.counters()
        if best is None or t["render"][0] < best[0]["render"][0]:
            best = (t, c)
    t, c = best
    tr = [ctx.trace_resident(kernel=0, any_hit=a, repeats=5) for a in (False, True)]
    print(json.dumps(dict(opts=opts, render_ms=round(t["render"][0], 2), mrays_s=round((c["extend"] + c["shadow"]) / t["render"][0] / 1e3, 1),
                          stages={k: round(v[0], 2) for k, v in t.items() if v[1]}, trace_closest_mrays=round(o.shape[0] / tr[0] / 1e3), trace_any_mrays=round(o.shape[0] / tr[1] / 1e3))))
