"""ncu target: one launch each of the one-ray-per-lane kernel (0) and the team kernel (2) on the same resident 2^21 incoherent rays.  python tools/team_ncu_target.py c3"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import path_tracer_b200 as ptb
from path_tracer_b200 import scenes

sd = {"c2": scenes.scene_c2, "c3": scenes.scene_c3, "c4": scenes.scene_c4}[sys.argv[1]]()
ctx = ptb.Context(0)
ctx.upload(sd)
o, d, tm = scenes.incoherent_rays(sd, 1 << 21)
ctx.trace_upload(o, d, tm)
for k in (0, 2):
    print(k, ctx.trace_resident(kernel=k, any_hit=False, repeats=1))
