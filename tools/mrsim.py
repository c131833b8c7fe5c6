"""SIMT cost model of the persistent traversal loop with K rays per lane (host emulation): python tools/mrsim.py c3
Calibration: K = 1 against profiles/r2m_c3_extend_sass_blocks.txt (stage instruction counts below are that file's)."""
import ctypes as C, os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "hostemu"))
so = os.path.join(ROOT, "gpurun_out", "libmrsim.so")
os.makedirs(os.path.dirname(so), exist_ok=True)
csrc = os.path.join(ROOT, "path_tracer_b200", "csrc")
subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-march=x86-64-v3", "-ffp-contract=off", "-pthread", "-I/usr/local/cuda/include", "-shared", "-o", so,
                       os.path.join(ROOT, "tools", "mrsim.cpp"), os.path.join(csrc, "host_build.cpp"), os.path.join(csrc, "host_wide.cpp")])
import emu
emu._LIB = C.CDLL(so); emu._LIB.emu_create.restype = C.c_void_p
from path_tracer_b200 import scenes
from path_tracer_b200.api import pack_scene
name = sys.argv[1] if len(sys.argv) > 1 else "c3"
sd = {"c2": scenes.scene_c2, "c3": scenes.scene_c3, "c4s": lambda: scenes.scene_c4(nverts=700)}[name]()
es = emu.EmuScene(sd, pack_scene)

def p(a): return a.ctypes.data_as(C.c_void_p)

# instruction counts of the stages (k_extend<7>, C3 capture): N 241, L 90, T 85, I 106, control per iteration 156, start+finish 1100
COST = dict(N=241, L=90, T=85, I=106, CTRL=156, START=550, FINISH=550)
# extra instructions per executed stage when ray state lives in shared memory (K > 1): loads/stores of the fields + ray selection
EXTRA = dict(N=30, L=30, T=25, I=35, CTRL=40)

def run(o, d, tm, any_hit, K, refill_below=8, quorum=8, policy=0, ql=8, qi=8, n_warps=256):
    n = o.shape[0]
    o, d, tm = (np.ascontiguousarray(x, np.float32) for x in (o, d, tm))
    out = np.zeros(17, np.uint64); hit = np.zeros(n, np.uint8); t = np.zeros(n, np.float32)
    emu._LIB.mrsim_run(C.c_void_p(es.h), 0, int(any_hit), C.c_uint64(n), p(o), p(d), p(tm), K, n_warps, refill_below, quorum, policy, ql, qi, p(out), p(hit), p(t))
    return out.astype(float), hit, t

def report(label, o, d, tm, any_hit, configs):
    n = o.shape[0]
    base = None
    for (K, rb, qu, pol, ql, qi) in configs:
        c, hit, t = run(o, d, tm, any_hit, K, rb, qu, pol, ql, qi)
        it = c[0]
        ex = EXTRA if K > 1 else dict.fromkeys(EXTRA, 0)
        cost = (it * (COST["CTRL"] + ex["CTRL"]) + c[1] * (COST["N"] + ex["N"]) + c[3] * (COST["L"] + ex["L"]) + c[5] * (COST["T"] + ex["T"]) + c[7] * (COST["I"] + ex["I"])
                + c[11] * COST["START"] + c[15] * COST["FINISH"])
        lanes_work = (c[2] * COST["N"] + c[4] * COST["L"] + c[6] * COST["T"] + c[8] * COST["I"])
        key = (hit.sum(), float(t.sum()))
        if base is None: base = (cost, key)
        assert key == base[1], (key, base[1])
        print(f"{label:18s} K={K} refill<{rb:2d} quorum={qu:2d} pol={pol} ql={ql} qi={qi}: warp-instr/ray {cost / n:7.1f} ({base[0] / cost:5.3f}x)  iters/ray {it * 32 / n:6.2f}  "
              f"N {c[1] / it:4.2f}@{c[2] / max(c[1], 1):5.2f}  L {c[3] / it:4.2f}@{c[4] / max(c[3], 1):5.2f}  T {c[5] / it:4.2f}@{c[6] / max(c[5], 1):5.2f}  "
              f"I {c[7] / it:4.2f}@{c[8] / max(c[7], 1):5.2f}  P @{c[10] / max(c[9], 1):5.2f}  rays/warp {c[13] / it:5.1f}  refills {c[11]:.0f}@{c[12] / max(c[11], 1):4.1f}")

W, H = 512, 288
st = scenes.pixel_centre_st(W, H)
co, cd = emu.camera_rays(sd.cam_origin, sd.cam_target, sd.fov_deg, sd.aspect, st)
r = es.trace(co, cd, np.full(co.shape[0], np.inf, np.float32))
m = r["hit"] > 0
P = co[m] + cd[m] * r["t"][m, None]
N = r["normal"][m]
rng = np.random.default_rng(1)
v = rng.normal(size=P.shape).astype(np.float32); v /= np.linalg.norm(v, axis=1, keepdims=True)
D = N + v * 0.999; D = (D / np.linalg.norm(D, axis=1, keepdims=True)).astype(np.float32)
P = (P + N * 1e-2).astype(np.float32)
inf = np.full(P.shape[0], np.inf, np.float32)
cfgs = [(1, 8, 8, 0, 0, 0), (1, 8, 8, 2, 8, 8), (2, 8, 8, 0, 0, 0), (2, 8, 16, 2, 8, 8), (2, 8, 16, 2, 12, 12), (3, 8, 16, 2, 12, 12), (4, 8, 16, 2, 12, 12), (4, 8, 16, 2, 16, 16), (4, 8, 20, 2, 20, 16), (6, 8, 20, 2, 20, 16), (8, 8, 24, 2, 24, 20)]
print("rays:", P.shape[0])
report("bounce closest", P, D, inf, False, cfgs)
report("primary closest", co, cd, np.full(co.shape[0], np.inf, np.float32), False, cfgs[:8])
tm = np.full(P.shape[0], 30.0, np.float32)
report("bounce any(t<30)", P, D, tm, True, cfgs[:8])
