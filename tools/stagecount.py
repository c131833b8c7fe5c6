"""Stage counts per ray of the 8-wide traversal (host emulation): python tools/stagecount.py c3"""
import ctypes as C, os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "hostemu"))
so = os.path.join(ROOT, "gpurun_out", "libstagecount.so")
os.makedirs(os.path.dirname(so), exist_ok=True)
csrc = os.path.join(ROOT, "path_tracer_b200", "csrc")
subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-march=x86-64-v3", "-ffp-contract=off", "-pthread", "-I/usr/local/cuda/include", "-shared", "-o", so,
                       os.path.join(ROOT, "tools", "stagecount.cpp"), os.path.join(csrc, "host_build.cpp"), os.path.join(csrc, "host_wide.cpp")])
import emu
emu._LIB = C.CDLL(so); emu._LIB.emu_create.restype = C.c_void_p
from path_tracer_b200 import scenes
from path_tracer_b200.api import pack_scene
name = sys.argv[1] if len(sys.argv) > 1 else "c3"
sd = {"c2": scenes.scene_c2, "c3": scenes.scene_c3, "c4s": lambda: scenes.scene_c4(nverts=700)}[name]()
es = emu.EmuScene(sd, pack_scene)
print("stats", es.stats())
def rd(reset=True):
    out = np.zeros(8, np.uint64); emu._LIB.emu_stat_read(out.ctypes.data_as(C.c_void_p), int(reset)); return out
names = ["blas_node", "leaf", "tri", "inst_box", "inst_enter", "tlas_node"]
def report(label, o, d, tm, any_hit):
    rd()
    r = es.trace(o, d, tm, any_hit=any_hit, threads=1)
    c = rd().astype(float) / o.shape[0]
    print(f"{label:28s} hit {r['hit'].mean():.3f} | " + "  ".join(f"{n} {c[i]:.2f}" for i, n in enumerate(names)))
    return r
n = 1 << 16
o, d, tm = scenes.incoherent_rays(sd, n)
report("incoherent closest", o, d, np.full(n, np.inf, np.float32), False)
report("incoherent any", o, d, tm, True)
W, H = 384, 216
st = scenes.pixel_centre_st(W, H)
co, cd = emu.camera_rays(sd.cam_origin, sd.cam_target, sd.fov_deg, sd.aspect, st)
r = report("primary closest", co, cd, np.full(co.shape[0], np.inf, np.float32), False)
# a diffuse bounce off the primary hits: cosine-ish random directions about the hit normal
m = r["hit"] > 0
P = co[m] + cd[m] * r["t"][m, None]
N = r["normal"][m]
rng = np.random.default_rng(1)
v = rng.normal(size=P.shape).astype(np.float32); v /= np.linalg.norm(v, axis=1, keepdims=True)
D = N + v * 0.999; D = (D / np.linalg.norm(D, axis=1, keepdims=True)).astype(np.float32)
P = (P + N * 1e-2).astype(np.float32)
report("bounce closest", P, D, np.full(P.shape[0], np.inf, np.float32), False)
