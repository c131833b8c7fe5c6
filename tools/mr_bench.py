"""Multiray traversal (kernels 3 / 4 / 5: K = 2 / 3 / 4 rays per lane with gated stages, kernel 6: K = 1 on shared-memory records) against
the one-ray-per-lane persistent kernel (kernel 0) on resident ray batches: parity and throughput.  python tools/mr_bench.py c3 [quick]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import path_tracer_b200 as ptb
from path_tracer_b200 import scenes

name = sys.argv[1]
quick = len(sys.argv) > 2 and sys.argv[2] == "quick"
sd = {"c2": scenes.scene_c2, "c3": scenes.scene_c3, "c4": scenes.scene_c4}[name]()
ctx = ptb.Context(0)
ctx.upload(sd)
ctx.set_camera(sd.cam_origin, sd.cam_target, sd.fov_deg, sd.aspect)
KN = {0: "lane", 6: "mr1", 3: "mr2", 4: "mr3", 5: "mr4"}


def opts(fq=24, qn=16, ql=12, qi=12, blocks=0):
    for k, v in (("mr_free_quorum", fq), ("mr_qn", qn), ("mr_ql", ql), ("mr_qi", qi), ("mr_blocks", blocks)):
        ctx.set_option(k, v)


def batches():
    n = 1 << 21
    o, d, tm = scenes.incoherent_rays(sd, n)
    yield "incoherent", o, d, tm
    st = scenes.pixel_centre_st(sd.width, sd.height, stride=max(1, int((sd.width * sd.height / n) ** 0.5)))[:n]
    po, pd = ctx.camera_rays(st)
    inf = np.full(po.shape[0], np.inf, np.float32)
    yield "primary", po, pd, inf
    # a diffuse bounce off the primary hits, in pixel order (what the wavefront's second iteration traces)
    r = ctx.trace_closest(po, pd, inf, kernel=0)
    m = r["hit"] > 0
    P = po[m] + pd[m] * r["t"][m, None]
    N = r["normal"][m]
    rng = np.random.default_rng(1)
    v = rng.normal(size=P.shape).astype(np.float32)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    D = N + v * np.float32(0.999)
    D = (D / np.linalg.norm(D, axis=1, keepdims=True)).astype(np.float32)
    P = (P + N * np.float32(1e-2)).astype(np.float32)
    reps = max(1, n // max(1, P.shape[0]))
    if reps > 1:
        P = np.repeat(P, reps, axis=0)[:n]
        D2 = np.repeat(D, reps, axis=0)[:n]
        v = rng.normal(size=P.shape).astype(np.float32) * np.float32(0.3)
        D = D2 + v
        D = (D / np.linalg.norm(D, axis=1, keepdims=True)).astype(np.float32)
    yield "bounce", P, D, np.full(P.shape[0], np.inf, np.float32)


first = True
for label, o, d, tm in batches():
    n = o.shape[0]
    # parity on a slice through the host path (downloads everything), against kernel 0
    if first or not quick:
        ns = min(n, 1 << 18)
        opts()
        ref = ctx.trace_closest(o[:ns], d[:ns], tm[:ns], kernel=0)
        tma = np.where(np.isfinite(tm[:ns]), tm[:ns], np.float32(30.0)).astype(np.float32)
        refa = ctx.trace_any(o[:ns], d[:ns], tma, kernel=0)
        for k in (6, 3, 4, 5):
            g = ctx.trace_closest(o[:ns], d[:ns], tm[:ns], kernel=k)
            ga = ctx.trace_any(o[:ns], d[:ns], tma, kernel=k)
            m = ref["hit"] > 0
            print(json.dumps(dict(scene=name, batch=label, parity=KN[k], rays=int(ns), hit_equal=bool(np.array_equal(g["hit"], ref["hit"])),
                                  t_bits_equal=bool(np.array_equal(g["t"][m].view(np.uint32), ref["t"][m].view(np.uint32))),
                                  uv_bits_equal=bool(np.array_equal(g["u"][m].view(np.uint32), ref["u"][m].view(np.uint32)) and np.array_equal(g["v"][m].view(np.uint32), ref["v"][m].view(np.uint32))),
                                  prim_diffs=int((g["prim"][m] != ref["prim"][m]).sum()), inst_diffs=int((g["inst"][m] != ref["inst"][m]).sum()),
                                  normal_equal=bool(np.array_equal(g["normal"][m], ref["normal"][m])), any_equal=bool(np.array_equal(ga["hit"], refa["hit"])))), flush=True)
    first = False
    ctx.trace_upload(o, d, tm)
    rec = dict(scene=name, batch=label, rays=int(n))
    for k in (0, 6, 3, 4, 5):
        opts()
        for a in (False, True):
            ctx.trace_resident(kernel=k, any_hit=a, repeats=2)
            ms = ctx.trace_resident(kernel=k, any_hit=a, repeats=6)
            rec[f"{KN[k]}_{'any' if a else 'closest'}_mrays"] = round(n / ms / 1e3, 1)
    print(json.dumps(rec), flush=True)
    if quick:
        continue
    # quorum / occupancy sweeps, closest hit
    for k in (6, 3, 4, 5):
        for (fq, qn, ql, qi, bl) in ((24, 16, 12, 12, 0), (24, 16, 8, 8, 0), (24, 8, 1, 1, 0), (24, 20, 16, 16, 0), (16, 16, 12, 12, 0), (40, 16, 12, 12, 0),
                                     (24, 16, 8, 8, 4), (24, 16, 8, 8, 5), (24, 16, 8, 8, 6), (24, 16, 8, 8, 7), (24, 16, 8, 8, 8)):
            if k == 6 and fq == 40:
                continue
            opts(fq, qn, ql, qi, bl)
            ctx.trace_resident(kernel=k, any_hit=False, repeats=1)
            ms = ctx.trace_resident(kernel=k, any_hit=False, repeats=4)
            print(json.dumps(dict(scene=name, batch=label, kernel=KN[k], fq=fq, qn=qn, ql=ql, qi=qi, blocks=bl, closest_mrays=round(n / ms / 1e3, 1))), flush=True)
