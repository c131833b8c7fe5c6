import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import path_tracer_b200 as ptb
from path_tracer_b200 import scenes
sd = scenes.scene_materials_test()
W, H = 96, 54
def run(dense, spp, tail, thr=0):
    ctx = ptb.Context(0); ctx.upload(sd)
    ctx.set_option("dense", dense); ctx.set_option("dense_threshold", thr); ctx.set_option("tail_threshold", tail)
    ctx.set_camera(sd.cam_origin, sd.cam_target, sd.fov_deg, np.float32(W) / np.float32(H))
    ctx.frame_begin(W, H); ctx.reset_counters()
    ctx.render(0, spp, seed=11, max_bounces=sd.max_bounces)
    pos, ids = ctx.read_aux()
    return ctx.read_accum(), pos, ids, ctx.counters()
for spp in (1, 4, 8):
    for tail in (-1, 0):
        a = run(0, spp, tail); b = run(1, spp, tail)
        da = (a[0].view(np.uint32) != b[0].view(np.uint32)).any(-1)
        print("spp", spp, "tail", tail, "accum px differ", int(da.sum()), "pos", int((a[1].view(np.uint32) != b[1].view(np.uint32)).any(-1).sum()), "ids", int((a[2] != b[2]).sum()),
              "w differ", int((a[0][..., 3] != b[0][..., 3]).sum()), {k: (a[3][k], b[3][k]) for k in ("extend", "shadow", "probe", "paths", "iterations")})
        if da.any():
            ys, xs = np.nonzero(da)
            print("   first", list(zip(ys[:5], xs[:5])), a[0][ys[0], xs[0]], b[0][ys[0], xs[0]])
