"""ctypes wrapper of the CPU oracle (oracle/libpt_oracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  Nothing under path_tracer_b200/ imports this module.  PARITY UNPINNED (see po_math.hpp).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class Material(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("colour", C.c_float * 3), ("roughness", C.c_float), ("ior", C.c_float), ("has_volume", C.c_uint32),
                ("vol_absorption", C.c_float * 3), ("vol_k", C.c_float), ("vol_c", C.c_float), ("vol_g", C.c_float)]


class BlasDesc(C.Structure):
    _fields_ = [("first_tri", C.c_uint32), ("tri_count", C.c_uint32), ("material", Material)]


class Instance(C.Structure):
    _fields_ = [("blas", C.c_uint32), ("affine", C.c_float * 12)]


def build(force=False):
    so = os.path.join(_HERE, "libpt_oracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("po_capi.cpp", "po_math.hpp", "po_accel.hpp", "po_shading.hpp", "po_temporal.hpp", "po_obj.hpp", "Makefile")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return so


BUILD_FLAGS = "-O2 -march=x86-64-v3 -ffp-contract=off (portable checker build)"


def use_native_build():
    """bench.py's CPU arms only: compile the oracle on THIS host with the reference's release settings (-O3, LTO, native CPU;
    oracle/Makefile `native`) and load that instead of the portable checker build.  Must be called before the first lib().
    Falls back to the portable build (and says so in BUILD_FLAGS) when the compiler is missing or the build fails."""
    global _LIB, BUILD_FLAGS
    if _LIB is not None:
        return BUILD_FLAGS
    so = os.path.join(_HERE, "_native", "libpt_oracle_native.so")
    try:
        subprocess.check_call(["make", "-C", _HERE, "-s", "native"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        _LIB = C.CDLL(so)
        _declare(_LIB)
        BUILD_FLAGS = "-O3 -march=native -ffp-contract=off, one translation unit, built on this host (reference: opt-level 3, thin LTO, target-cpu=native; Cargo.toml:10-11, .cargo/config.toml:1-2)"
    except Exception as e:  # noqa: BLE001
        _LIB = None
        BUILD_FLAGS += f"; native build unavailable: {type(e).__name__}"
    return BUILD_FLAGS


def _declare(L):
    L.po_last_error.restype = C.c_char_p
    L.po_scene_create.restype = C.c_void_p
    L.po_scene_create.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32]
    L.po_scene_destroy.argtypes = [C.c_void_p]
    L.po_render.restype = C.c_double
    L.po_generate_blas_span.restype = C.c_uint32
    L.po_light_count.restype = C.c_uint32
    L.po_light_count.argtypes = [C.c_void_p]
    L.po_default_threads.restype = C.c_int
    for f in ("po_lk_hash", "po_low_bias_hash", "po_scramble_base2"):
        getattr(L, f).restype = C.c_uint32


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _declare(_LIB)
        return _LIB
    return _LIB


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def pack_scene(sd):
    """SceneData -> (verts f32 [ntri,3,6], BlasDesc array, Instance array)"""
    verts = np.ascontiguousarray(sd.verts, dtype=np.float32)
    bd = (BlasDesc * len(sd.blas))()
    for i, (first, cnt, m) in enumerate(sd.blas):
        bd[i].first_tri, bd[i].tri_count = first, cnt
        mm = bd[i].material
        mm.kind = m["kind"]
        mm.colour[:] = m["colour"]
        mm.roughness, mm.ior = m["roughness"], m["ior"]
        if m.get("volume"):
            a, k, c, g = m["volume"]
            mm.has_volume = 1
            mm.vol_absorption[:] = a
            mm.vol_k, mm.vol_c, mm.vol_g = k, c, g
    inst = (Instance * len(sd.instances))()
    for i, (b, aff) in enumerate(sd.instances):
        inst[i].blas = b
        inst[i].affine[:] = aff
    return verts, bd, inst


class OracleScene:
    """Scene::new (scene.rs:21-35) over flat arrays, plus the probes the parity tests need."""

    def __init__(self, sd):
        L = lib()
        self.sd = sd
        verts, bd, inst = pack_scene(sd)
        self._keep = (verts, bd, inst)
        self.h = L.po_scene_create(_p(verts), C.c_uint64(verts.shape[0]), C.byref(bd), len(sd.blas), C.byref(inst), len(sd.instances))
        if not self.h:
            raise ValueError(L.po_last_error().decode())
        self.nblas = len(sd.blas)

    def set_env(self, rgb):
        """Environment map, linear RGB [H, W, 3] (ImageHelper::load_image output, image_helper.rs:25-33)."""
        rgb = np.ascontiguousarray(rgb, np.float32)
        if lib().po_scene_set_env(self.h, rgb.shape[1], rgb.shape[0], _p(rgb)) != 0:
            raise RuntimeError(lib().po_last_error().decode())

    def close(self):
        if self.h:
            lib().po_scene_destroy(C.c_void_p(self.h))
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- canonical dump ---------------------------------------------------------------------------
    def blas_dump(self, b):
        L = lib()
        nc, lc, root = C.c_uint32(), C.c_uint32(), C.c_uint32()
        L.po_blas_counts(C.c_void_p(self.h), b, C.byref(nc), C.byref(lc), C.byref(root))
        boxes = np.zeros((nc.value, 6), np.float32)
        meta = np.zeros((nc.value, 3), np.uint32)
        lp = np.zeros(max(lc.value, 1), np.uint32)
        L.po_blas_nodes(C.c_void_p(self.h), b, _p(boxes), _p(meta), _p(lp))
        tri = np.zeros((self.sd.blas[b][1], 30), np.float32)
        L.po_blas_triangles(C.c_void_p(self.h), b, _p(tri))
        return dict(boxes=boxes, meta=meta, leaf_prims=lp[: lc.value], root=root.value, triangles=tri)

    def tlas_dump(self, which=0):
        L = lib()
        nc, root, nb = C.c_uint32(), C.c_uint32(), C.c_uint32()
        L.po_tlas_counts(C.c_void_p(self.h), which, C.byref(nc), C.byref(root), C.byref(nb))
        boxes = np.zeros((nc.value, 6), np.float32)
        meta = np.zeros((nc.value, 3), np.uint32)
        mats = np.zeros((nc.value, 24), np.float32)
        ids = np.zeros(nb.value, np.uint32)
        L.po_tlas_nodes(C.c_void_p(self.h), which, _p(boxes), _p(meta), _p(mats), _p(ids))
        return dict(boxes=boxes, meta=meta, matrices=mats, blas_ids=ids, root=root.value)

    def light_sampler(self):
        L = lib()
        n = L.po_light_count(C.c_void_p(self.h))
        items = np.zeros((n, 2), np.uint32)
        pdf = np.zeros(n, np.float32)
        cdf = np.zeros(n, np.float32)
        mx = C.c_float()
        L.po_light_sampler(C.c_void_p(self.h), _p(items), _p(pdf), _p(cdf), C.byref(mx))
        return dict(items=items, pdf=pdf, cdf=cdf, max=np.float32(mx.value))

    # ---- traversal --------------------------------------------------------------------------------
    def trace_closest(self, o, d, tmax, which=0, threads=0):
        n = o.shape[0]
        o, d, tmax = (np.ascontiguousarray(x, np.float32) for x in (o, d, tmax))
        out = dict(hit=np.zeros(n, np.uint8), t=np.zeros(n, np.float32), u=np.zeros(n, np.float32), v=np.zeros(n, np.float32), prim=np.zeros(n, np.uint32),
                   blas=np.zeros(n, np.uint32), inst=np.zeros(n, np.uint32), front=np.zeros(n, np.uint8), normal=np.zeros((n, 3), np.float32))
        cnt = np.zeros(3, np.uint64)
        lib().po_trace_closest(C.c_void_p(self.h), which, C.c_uint64(n), _p(o), _p(d), _p(tmax), _p(out["hit"]), _p(out["t"]), _p(out["u"]), _p(out["v"]), _p(out["prim"]),
                               _p(out["blas"]), _p(out["inst"]), _p(out["front"]), _p(out["normal"]), _p(cnt), threads)
        out["counters"] = dict(boxes=int(cnt[0]), tris=int(cnt[1]), insts=int(cnt[2]))
        return out

    def trace_any(self, o, d, tmax, which=0, threads=0):
        n = o.shape[0]
        o, d, tmax = (np.ascontiguousarray(x, np.float32) for x in (o, d, tmax))
        hit = np.zeros(n, np.uint8)
        cnt = np.zeros(3, np.uint64)
        lib().po_trace_any(C.c_void_p(self.h), which, C.c_uint64(n), _p(o), _p(d), _p(tmax), _p(hit), _p(cnt), threads)
        return dict(hit=hit, counters=dict(boxes=int(cnt[0]), tris=int(cnt[1]), insts=int(cnt[2])))

    def trace_brute(self, o, d, tmax, which=0, threads=0):
        n = o.shape[0]
        o, d, tmax = (np.ascontiguousarray(x, np.float32) for x in (o, d, tmax))
        out = dict(hit=np.zeros(n, np.uint8), t=np.zeros(n, np.float32), prim=np.zeros(n, np.uint32), inst=np.zeros(n, np.uint32))
        lib().po_trace_brute_closest(C.c_void_p(self.h), which, C.c_uint64(n), _p(o), _p(d), _p(tmax), _p(out["hit"]), _p(out["t"]), _p(out["prim"]), _p(out["inst"]), threads)
        return out

    def eval_candidate(self, inst, prim, o3, d3, tmax, which=0):
        o3 = np.ascontiguousarray(o3, np.float32)
        d3 = np.ascontiguousarray(d3, np.float32)
        t = C.c_float()
        ok = lib().po_eval_candidate(C.c_void_p(self.h), which, int(inst), int(prim), _p(o3), _p(d3), C.c_float(float(tmax)), C.byref(t))
        return bool(ok), np.float32(t.value)

    # ---- frame driver -----------------------------------------------------------------------------
    def render(self, width=None, height=None, sample_begin=0, sample_count=1, seed=1, max_bounces=None, enable_nee=True, threads=0, accum=None, want_sq=False,
               cam=None):
        sd = self.sd
        W, H = width or sd.width, height or sd.height
        mb = sd.max_bounces if max_bounces is None else max_bounces
        if accum is None:
            accum = np.zeros((H, W, 4), np.float32)
        sq = np.zeros((H, W, 4), np.float32) if want_sq else None
        pos = np.zeros((H, W, 4), np.float32)
        ids = np.zeros((H, W), np.uint32)
        cnt = np.zeros(14, np.uint64)
        org = np.asarray(cam[0] if cam else sd.cam_origin, np.float32)
        tgt = np.asarray(cam[1] if cam else sd.cam_target, np.float32)
        aspect = np.float32(W) / np.float32(H)
        secs = lib().po_render(C.c_void_p(self.h), _p(org), _p(tgt), C.c_float(sd.fov_deg), C.c_float(aspect), W, H, sample_begin, sample_count, C.c_uint64(seed), mb,
                               int(enable_nee), threads, _p(accum), _p(sq), _p(pos), _p(ids), _p(cnt))
        names = ["extend", "shadow", "probe", "paths", "bounces", "extend_boxes", "extend_tris", "extend_insts", "light_boxes", "light_tris", "light_insts", "shadow_boxes", "shadow_tris", "shadow_insts"]
        return dict(accum=accum, accum_sq=sq, position=pos, id=ids, seconds=secs, counters={k: int(v) for k, v in zip(names, cnt)})


def load_obj_text(text):
    """load_obj (blas.rs:44-131) on a text image -> float32 [ntri, 3, 6]; ValueError where the reference panics."""
    b = text.encode() if isinstance(text, str) else bytes(text)
    n = C.c_uint64()
    L = lib()
    if L.po_load_obj_text(b, C.c_uint64(len(b)), None, C.byref(n)):
        raise ValueError(L.po_last_error().decode())
    out = np.zeros((n.value, 3, 6), np.float32)
    if n.value:
        L.po_load_obj_text(b, C.c_uint64(len(b)), _p(out), C.byref(n))
    return out


def generate_blas_span(boxes, prims, last_axis=4):
    """Re-run the restated generate_blas (blas_bvh.rs:62-136) on an ordered span."""
    L = lib()
    boxes = np.ascontiguousarray(boxes, np.float32)
    prims = np.ascontiguousarray(prims, np.uint32)
    n = boxes.shape[0]
    lc = C.c_uint32()
    nc = L.po_generate_blas_span(_p(boxes), _p(prims), C.c_uint64(n), last_axis, None, None, None, C.byref(lc))
    ob = np.zeros((nc, 6), np.float32)
    om = np.zeros((nc, 3), np.uint32)
    ol = np.zeros(max(lc.value, 1), np.uint32)
    L.po_generate_blas_span(_p(boxes), _p(prims), C.c_uint64(n), last_axis, _p(ob), _p(om), _p(ol), C.byref(lc))
    return dict(boxes=ob, meta=om, leaf_prims=ol[: lc.value], root=nc - 1)


def sobol_table():
    t = np.zeros((512, 2), np.uint32)
    lib().po_sobol_table(_p(t))
    return t


def ss_sobol(index, seed):
    index = np.ascontiguousarray(index, np.uint32)
    seed = np.ascontiguousarray(seed, np.uint32)
    out = np.zeros((index.shape[0], 2), np.float32)
    lib().po_ss_sobol(C.c_uint64(index.shape[0]), _p(index), _p(seed), _p(out))
    return out


def camera_rays(origin, target, fov_deg, aspect, st):
    st = np.ascontiguousarray(st, np.float32)
    n = st.shape[0]
    o = np.zeros((n, 3), np.float32)
    d = np.zeros((n, 3), np.float32)
    org = np.asarray(origin, np.float32)
    tgt = np.asarray(target, np.float32)
    lib().po_camera_rays(_p(org), _p(tgt), C.c_float(fov_deg), C.c_float(aspect), C.c_uint64(n), _p(st), _p(o), _p(d))
    return o, d


def rng_stream(seed, pixel, sample, n):
    out = np.zeros(n, np.uint32)
    lib().po_rng_stream(C.c_uint64(seed), pixel, sample, n, _p(out))
    return out


def tonemap_rgb8(accum, flip_y=True):
    """ImageHelper::write_image's conversion (image_helper.rs:44-46) of an accumulation buffer [H,W,4] (sum rgb, count)."""
    accum = np.ascontiguousarray(accum, np.float32)
    H, W = accum.shape[:2]
    out = np.zeros((H, W, 3), np.uint8)
    lib().po_tonemap_rgb8(_p(accum), W, H, int(flip_y), _p(out))
    return out


def state_update(data, position, ids, inv_projection, last_inv_projection, accumulation):
    """State::update (state.rs:505-579) on host arrays.  data/position/accumulation [H,W,4] f32, ids [H,W] u32, matrices 16 floats
    column-major.  Returns (new accumulation, velocity [H,W,2], taa_ran)."""
    data = np.ascontiguousarray(data, np.float32)
    position = np.ascontiguousarray(position, np.float32)
    ids = np.ascontiguousarray(ids, np.uint32)
    acc = np.array(accumulation, np.float32, order="C", copy=True)
    H, W = data.shape[:2]
    vel = np.zeros((H, W, 2), np.float32)
    a = np.ascontiguousarray(np.asarray(inv_projection, np.float32).reshape(16))
    b = np.ascontiguousarray(np.asarray(last_inv_projection, np.float32).reshape(16))
    f = lib().po_state_update
    f.restype = C.c_int
    ran = f(C.c_uint32(W), C.c_uint32(H), _p(data), _p(position), _p(ids), _p(a), _p(b), _p(acc), _p(vel))
    return acc, vel, bool(ran)


def view_projection(origin, target, fov_deg, aspect):
    """World -> clip matrix the reference passes as `inv_projection` (main.rs:209-216), float64, column-major [16]."""
    o = np.ascontiguousarray(origin, np.float32)
    t = np.ascontiguousarray(target, np.float32)
    m = np.zeros(16, np.float64)
    f = lib().po_view_projection
    f.restype = None
    f(_p(o), _p(t), C.c_float(fov_deg), C.c_float(aspect), _p(m))
    return m
