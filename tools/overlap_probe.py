"""Probe: two contexts on ONE GPU rendering alternate frames (ptb_render_async / ptb_wait): does one frame's thin phase / tail
overlap the other's dense phase?  python tools/overlap_probe.py c3 64 12 [opt=value ...]   (frames in total)"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import path_tracer_b200 as ptb
from path_tracer_b200 import scenes

name, spp, frames = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
opts = [a.split("=") for a in sys.argv[4:]]
sd = {"c2": scenes.scene_c2, "c3": scenes.scene_c3, "c4": scenes.scene_c4}[name]()


def make():
    c = ptb.Context(0)
    c.upload(sd)
    c.set_camera(sd.cam_origin, sd.cam_target, sd.fov_deg, sd.aspect)
    c.frame_begin(sd.width, sd.height)
    c.render(0, 2, seed=1, max_bounces=sd.max_bounces)
    return c


ctxs = [make(), make()]
out = np.zeros((sd.height, sd.width, 4), np.float32)


def serial(n):
    c = ctxs[0]
    t = time.perf_counter()
    for i in range(n):
        c.frame_begin(sd.width, sd.height)
        c.render(i * spp, spp, seed=1, max_bounces=sd.max_bounces)
        c.read_accum(out)
    return (time.perf_counter() - t) / n


def pipelined(n):
    t = time.perf_counter()
    for i in range(n + 2):
        c = ctxs[i % 2]
        if i >= 2:
            c.wait()
            c.read_accum(out)
        if i < n:
            c.frame_begin(sd.width, sd.height)
            c.render_async(i * spp, spp, seed=1, max_bounces=sd.max_bounces)
    return (time.perf_counter() - t) / n


base = serial(frames // 2)
print(f"{name} {spp} spp: one context, frame after frame: {base * 1e3:.1f} ms/frame (incl. read-back)")
variants = [dict(), dict(shared_device=1), dict(shared_device=1, thin_rays_per_lane=2), dict(shared_device=1, thin_rays_per_lane=4), dict(shared_device=1, thin_rays_per_lane=8),
            dict(thin_rays_per_lane=4)]
for v in variants:
    for c in ctxs:
        for k in ("shared_device", "thin_rays_per_lane"):
            c.set_option(k, v.get(k, 0))
        for k, x in opts:
            c.set_option(k, int(x))
    pipelined(4)
    p = pipelined(frames)
    print(f"  two contexts alternating {v} {opts}: {p * 1e3:.1f} ms/frame  ({base / p:.3f} x)")
for c in ctxs:
    c.set_option("shared_device", 0)
    c.set_option("thin_rays_per_lane", 4)
print(f"  one context with thin_rays_per_lane=4: {serial(frames // 2) * 1e3:.1f} ms/frame")
