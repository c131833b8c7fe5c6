"""wavefront_groups must not change results: bit-identical accumulators, aux outputs and ray counts (tuning aid; the test lives in tests/test_gpu_parity.py)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import path_tracer_b200 as ptb
from path_tracer_b200 import scenes
for name, W, H, spp in (("c3", 320, 180, 20), ("c2", 333, 187, 33), ("c1", 160, 90, 8)):
    sd = {"c1": scenes.scene_c1, "c2": scenes.scene_c2, "c3": scenes.scene_c3}[name]()
    ctx = ptb.Context(0); ctx.upload(sd)
    ctx.set_camera(sd.cam_origin, sd.cam_target, sd.fov_deg, np.float32(W) / np.float32(H))
    res = {}
    for G in (1, 2, 3, 8):
        ctx.set_option("wavefront_groups", G)
        ctx.frame_begin(W, H); ctx.reset_counters()
        ctx.render(0, spp, seed=5, max_bounces=sd.max_bounces)
        pos, ids = ctx.read_aux()
        res[G] = (ctx.read_accum().copy(), pos.copy(), ids.copy(), ctx.counters())
    for G in (2, 3, 8):
        same = all(np.array_equal(res[1][k].view(np.uint32), res[G][k].view(np.uint32)) for k in range(3))
        c1, cg = res[1][3], res[G][3]
        print(name, "G", G, "bit-identical", same, "counters equal", all(c1[k] == cg[k] for k in ("extend", "shadow", "probe", "paths")), cg)
