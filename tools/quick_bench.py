"""Quick device-side throughput probe (not the contract bench): Mrays/s and per-stage time split for a scene."""
import sys, os, time, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import path_tracer_b200 as ptb
from path_tracer_b200 import api as _api
if os.environ.get("PTB_LIB"):  # A/B runs of two builds in one gpurun call (tuning aid)
    _api.lib_path = lambda: os.environ["PTB_LIB"]
from path_tracer_b200 import scenes

name = sys.argv[1] if len(sys.argv) > 1 else "c2"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 16
opts = dict(a.split("=") for a in sys.argv[3:])
sd = {"c1": scenes.scene_c1, "c2": scenes.scene_c2, "c3": scenes.scene_c3, "c4": scenes.scene_c4, "c4s": lambda: scenes.scene_c4(nverts=700)}[name]()
t0 = time.time(); ctx = ptb.Context(0); ctx.upload(sd); t_build = time.time() - t0
for k, v in opts.items():
    ctx.set_option(k, int(v))
W, H = sd.width, sd.height
ctx.set_camera(sd.cam_origin, sd.cam_target, sd.fov_deg, sd.aspect)
ctx.frame_begin(W, H)
ctx.render(0, 2, seed=1, max_bounces=sd.max_bounces)  # warm-up
ctx.frame_begin(W, H); ctx.reset_counters(); ctx.set_profiling(True)
t0 = time.time(); ctx.render(0, spp, seed=1, max_bounces=sd.max_bounces); ctx.synchronize(); wall = time.time() - t0
c = ctx.counters(); t = ctx.timing()
rays = c["extend"] + c["shadow"]
ms = t["render"][0]
log = ctx.iteration_log()
print("iterations", len(log), "active@", {int(i): int(log[i]) for i in sorted(set([0, 1, 2, 4, 8, 16, 32, 64, 128, 256, 512, len(log) - 1])) if i < len(log)})
print(json.dumps(dict(scene=sd.name, W=W, H=H, spp=spp, build_s=round(t_build, 2), stats=ctx.bvh_stats(), counters=c, render_ms=ms, wall_ms=wall * 1e3,
                      mrays_s=rays / ms / 1e3, timing={k: (round(v[0], 2), v[1]) for k, v in t.items()}, opts=opts)))
# trace-only probes
for label, (o, d, tm) in (("incoherent", scenes.incoherent_rays(sd, 1 << 21)),):
    ctx.trace_upload(o, d, tm)
    for kern in (0, 1):
        for anyh in (False, True):
            ctx.trace_resident(kernel=kern, any_hit=anyh, repeats=1)
            ms = ctx.trace_resident(kernel=kern, any_hit=anyh, repeats=5)
            print(f"trace {label} kernel={kern} any={anyh}: {ms:.3f} ms  {o.shape[0] / ms / 1e3:.1f} Mrays/s")
