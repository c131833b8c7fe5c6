"""Active slots per wavefront iteration + stage times (tuning aid): python tools/iterlog.py c3 64 [opt=value ...]"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import path_tracer_b200 as ptb
from path_tracer_b200 import api as _api
if os.environ.get("PTB_LIB"):
    _api.lib_path = lambda: os.environ["PTB_LIB"]
from path_tracer_b200 import scenes
name, spp = sys.argv[1], int(sys.argv[2])
sd = {"c2": scenes.scene_c2, "c3": scenes.scene_c3, "c4": scenes.scene_c4}[name]()
ctx = ptb.Context(0); ctx.upload(sd)
for a in sys.argv[3:]:
    k, v = a.split("="); ctx.set_option(k, int(v))
ctx.set_camera(sd.cam_origin, sd.cam_target, sd.fov_deg, sd.aspect)
ctx.frame_begin(sd.width, sd.height); ctx.render(0, 2, seed=1, max_bounces=sd.max_bounces)
ctx.frame_begin(sd.width, sd.height); ctx.reset_counters(); ctx.set_profiling(True)
ctx.render(0, spp, seed=1, max_bounces=sd.max_bounces); ctx.synchronize()
log = ctx.iteration_log()
print("iters", len(log)); print("active", [int(x) for x in log])
print("timing", {k: (round(v[0], 2), v[1]) for k, v in ctx.timing().items()}); print("counters", ctx.counters())
