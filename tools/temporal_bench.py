"""Times the temporal-pass kernels (k_resolve, k_accumulate; temporal.cu) on device-resident frame buffers and reports them against the
HBM roofline.  Tuning aid, not the contract bench: python tools/temporal_bench.py [1920 1080] [reps]
Algorithmic bytes per texel: k_resolve 16 input + 16 position + 8 ids + 16 history + 16 out + 8 velocity = 80 B;
k_accumulate 16 + 16 + 16 = 48 B."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import path_tracer_b200 as ptb
from path_tracer_b200 import scenes

W, H = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1920, 1080)
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 50
peak = 6555.2
try:
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
sd = scenes.scene_c2()
ctx = ptb.Context(0)
ctx.upload(sd)
cams = [(sd.cam_origin, sd.cam_target), ((sd.cam_origin[0] + 15.0, sd.cam_origin[1] + 4.0, sd.cam_origin[2]), sd.cam_target)]
ctx.set_camera(*cams[0], sd.fov_deg, sd.aspect)
vps = []
for c in cams:
    ctx.set_camera(*c, sd.fov_deg, sd.aspect)
    vps.append(ctx.view_projection())
ctx.frame_begin(W, H)
ctx.state_reset(W, H)
# two frames so that the id history is populated, then the timed updates re-run on the same resident buffers
for f in range(2):
    if f:
        ctx.frame_next()
    ctx.set_camera(*cams[f], sd.fov_deg, sd.aspect)
    ctx.render(f, 1, seed=1, max_bounces=sd.max_bounces)
    ctx.state_update_resident(vps[f], vps[0])
out = {"width": W, "height": H, "reps": reps, "hbm_peak_gbs": peak}
for name, (a, b), nbytes in (("k_resolve", (vps[1], vps[0]), 80), ("k_accumulate", (vps[1], vps[1]), 48)):
    for _ in range(5):
        ctx.state_update_resident(a, b)
    ctx.set_profiling(True)
    ctx.state_timing(reset=True)
    for _ in range(reps):
        ctx.state_update_resident(a, b)
    ms, n = ctx.state_timing(reset=True)
    ctx.set_profiling(False)
    us = ms / n * 1e3
    gbs = W * H * nbytes / (us * 1e-6) / 1e9
    out[name] = {"us": round(us, 2), "alg_bytes_per_texel": nbytes, "gb_s": round(gbs, 1), "frac_of_hbm_peak": round(gbs / peak, 3)}
print(json.dumps(out))
