"""ctypes binding of libptb.so (include/ptb.h) and the host-side mirror of the reference's render interface.

The reference is a Rust binary without a plugin API; its seam is `Scene::new(Vec<Model>)` (scene.rs:21),
`Camera::new` (camera.rs:17) and the per-frame pixel loop calling `integrator::integrate` (main.rs:181-207).
`Scene`, `Camera`, `Model` and the material constructors below keep those names and argument meanings; every
compute call goes through the C ABI to the CUDA kernels.  There is no CPU fallback: if libptb.so is missing or no
sm_100 device is present the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import scenes as _scenes

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

PTB_OK, PTB_ERR_INVALID, PTB_ERR_CUDA, PTB_ERR_NO_DEVICE, PTB_ERR_OOM, PTB_ERR_NOT_BUILT, PTB_ERR_UNSUPPORTED, PTB_ERR_NCCL = range(8)
_STATUS = {1: "PTB_ERR_INVALID", 2: "PTB_ERR_CUDA", 3: "PTB_ERR_NO_DEVICE", 4: "PTB_ERR_OOM", 5: "PTB_ERR_NOT_BUILT", 6: "PTB_ERR_UNSUPPORTED", 7: "PTB_ERR_NCCL"}

EXPORTS = [
    "ptb_abi_version", "ptb_create", "ptb_destroy", "ptb_last_error", "ptb_synchronize", "ptb_scene_upload", "ptb_scene_set_env", "ptb_build",
    "ptb_set_build_threads", "ptb_bvh_blas_counts", "ptb_bvh_blas_nodes", "ptb_bvh_blas_triangles", "ptb_bvh_tlas_counts", "ptb_bvh_tlas_nodes",
    "ptb_light_count", "ptb_light_sampler", "ptb_bvh_stats", "ptb_trace_closest", "ptb_trace_any", "ptb_trace_upload", "ptb_trace_resident",
    "ptb_trace_download", "ptb_set_camera", "ptb_camera_rays", "ptb_ss_sobol", "ptb_rng_stream", "ptb_frame_begin", "ptb_render", "ptb_render_async", "ptb_wait", "ptb_read_accum",
    "ptb_read_aux", "ptb_read_rgb8", "ptb_enable_second_moment", "ptb_read_second_moment", "ptb_accum_device_ptr", "ptb_bind_accum", "ptb_get_counters",
    "ptb_reset_counters", "ptb_set_profiling", "ptb_get_timing", "ptb_get_iteration_log", "ptb_get_launch_log", "ptb_set_option",
    "ptb_state_reset", "ptb_state_update", "ptb_state_update_resident", "ptb_state_read", "ptb_state_read_rgb8", "ptb_state_device_ptr", "ptb_state_timing", "ptb_frame_next",
    "ptb_camera_view_projection",
    "ptb_group_create", "ptb_group_unique_id", "ptb_group_join", "ptb_group_destroy", "ptb_group_last_error", "ptb_group_size", "ptb_group_context",
    "ptb_group_sample_range", "ptb_group_scene_upload", "ptb_group_build", "ptb_group_set_camera", "ptb_group_set_option", "ptb_group_frame_begin",
    "ptb_group_render", "ptb_group_reduce", "ptb_group_read_accum", "ptb_group_read_aux", "ptb_group_get_counters", "ptb_group_reset_counters",
    "ptb_group_reduce_ms", "ptb_load_obj", "ptb_load_obj_text", "ptb_load_obj_error", "ptb_free", "ptb_trace_closest_static",
]
STATIC_TRACE_MAX_RAYS = 65536  # PTB_STATIC_TRACE_MAX_RAYS
# ptb_get_counters: [0] extend rays that ran a traversal, [7] primary rays settled by the bounds pre-test without one (k_extend)
COUNTER_NAMES = ["extend", "shadow", "probe", "paths", "iterations", "launches", "extend_path", "settled"]


class PtbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"{_STATUS.get(code, code)}: {msg}")
        self.code = code


class Material(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("colour", C.c_float * 3), ("roughness", C.c_float), ("ior", C.c_float), ("has_volume", C.c_uint32),
                ("vol_absorption", C.c_float * 3), ("vol_k", C.c_float), ("vol_c", C.c_float), ("vol_g", C.c_float)]


class BlasDesc(C.Structure):
    _fields_ = [("first_tri", C.c_uint32), ("tri_count", C.c_uint32), ("material", Material)]


class Instance(C.Structure):
    _fields_ = [("blas", C.c_uint32), ("affine", C.c_float * 12)]


def lib_path():
    return os.path.join(_HERE, "libptb.so")


def _preload_nccl():
    """libptb.so links libnccl.so.2 (ptb_group_*).  PyTorch ships its own, newer libnccl.so.2 under the same SONAME: whichever is
    loaded first serves both, and torch fails to import on top of the older system library.  So in a Python process the copy
    bundled with torch is loaded first when there is one; a host without Python resolves the system library as usual."""
    import importlib.util
    try:
        spec = importlib.util.find_spec("nvidia")
    except (ImportError, ValueError):
        spec = None
    for base in (spec.submodule_search_locations if spec and spec.submodule_search_locations else []):
        cand = os.path.join(base, "nccl", "lib", "libnccl.so.2")
        if os.path.exists(cand):
            try:
                C.CDLL(cand, mode=C.RTLD_GLOBAL)
                return cand
            except OSError:
                pass
    return None


def lib():
    """Load libptb.so.  Fails loudly when the CUDA extension has not been built (no fallback)."""
    global _LIB
    if _LIB is None:
        p = lib_path()
        if not os.path.exists(p):
            raise ImportError(f"{p} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` (make -C path_tracer_b200/csrc)")
        _preload_nccl()
        L = C.CDLL(p)
        L.ptb_last_error.restype = C.c_char_p
        L.ptb_last_error.argtypes = [C.c_void_p]
        L.ptb_destroy.restype = None
        L.ptb_destroy.argtypes = [C.c_void_p]
        _LIB = L
    return _LIB


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def pack_scene(sd):
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


class Context:
    """One ptb_context: one GPU (device >= 0) or a host-only context (device = -1) that can only build and dump."""

    def __init__(self, device=0):
        self.L = lib()
        h = C.c_void_p()
        rc = self.L.ptb_create(int(device), C.byref(h))
        if rc != PTB_OK:
            raise PtbError(rc, self.L.ptb_last_error(None).decode())
        self.h = h
        self.device = device
        self.width = self.height = 0
        self._owned = True

    @classmethod
    def borrow(cls, handle, device):
        """View of a ptb_context owned by someone else (a member of a ptb_group): never destroyed from here."""
        self = cls.__new__(cls)
        self.L = lib()
        self.h = C.c_void_p(handle)
        self.device = device
        self.width = self.height = 0
        self._owned = False
        return self

    def close(self):
        if getattr(self, "h", None):
            if self._owned:
                self.L.ptb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != PTB_OK:
            raise PtbError(rc, self.L.ptb_last_error(self.h).decode())

    # ---- Scene::new ---------------------------------------------------------------------------------
    def upload(self, sd, build=True, threads=0):
        verts, bd, inst = pack_scene(sd)
        self._ck(self.L.ptb_scene_upload(self.h, _p(verts), C.c_uint64(verts.shape[0]), C.byref(bd), len(sd.blas), C.byref(inst), len(sd.instances)))
        self.sd = sd
        if build:
            self.build(threads)

    def build(self, threads=0):
        self._ck(self.L.ptb_set_build_threads(self.h, threads))
        self._ck(self.L.ptb_build(self.h))

    def set_env(self, rgb):
        rgb = np.ascontiguousarray(rgb, np.float32)
        self._ck(self.L.ptb_scene_set_env(self.h, rgb.shape[1], rgb.shape[0], _p(rgb)))

    def blas_dump(self, b):
        nc, lc, root = C.c_uint32(), C.c_uint32(), C.c_uint32()
        self._ck(self.L.ptb_bvh_blas_counts(self.h, b, C.byref(nc), C.byref(lc), C.byref(root)))
        boxes = np.zeros((nc.value, 6), np.float32)
        meta = np.zeros((nc.value, 3), np.uint32)
        lp = np.zeros(max(lc.value, 1), np.uint32)
        self._ck(self.L.ptb_bvh_blas_nodes(self.h, b, _p(boxes), _p(meta), _p(lp)))
        tri = np.zeros((self.sd.blas[b][1], 30), np.float32)
        self._ck(self.L.ptb_bvh_blas_triangles(self.h, b, _p(tri)))
        return dict(boxes=boxes, meta=meta, leaf_prims=lp[: lc.value], root=root.value, triangles=tri)

    def tlas_dump(self, which=0):
        nc, root, nb = C.c_uint32(), C.c_uint32(), C.c_uint32()
        self._ck(self.L.ptb_bvh_tlas_counts(self.h, which, C.byref(nc), C.byref(root), C.byref(nb)))
        boxes = np.zeros((nc.value, 6), np.float32)
        meta = np.zeros((nc.value, 3), np.uint32)
        mats = np.zeros((nc.value, 24), np.float32)
        ids = np.zeros(nb.value, np.uint32)
        self._ck(self.L.ptb_bvh_tlas_nodes(self.h, which, _p(boxes), _p(meta), _p(mats), _p(ids)))
        return dict(boxes=boxes, meta=meta, matrices=mats, blas_ids=ids, root=root.value)

    def light_sampler(self):
        n = C.c_uint32()
        self._ck(self.L.ptb_light_count(self.h, C.byref(n)))
        items = np.zeros((n.value, 2), np.uint32)
        pdf = np.zeros(n.value, np.float32)
        cdf = np.zeros(n.value, np.float32)
        mx = C.c_float()
        self._ck(self.L.ptb_light_sampler(self.h, _p(items), _p(pdf), _p(cdf), C.byref(mx)))
        return dict(items=items, pdf=pdf, cdf=cdf, max=np.float32(mx.value))

    def bvh_stats(self):
        out = np.zeros(8, np.uint64)
        self._ck(self.L.ptb_bvh_stats(self.h, _p(out)))
        names = ["wide_nodes", "leaf_headers", "tri_slots", "wide_bytes", "wide_depth", "bin_nodes", "instances", "stack_bound"]
        return {k: int(v) for k, v in zip(names, out)}

    # ---- TLAS::intersect / any_intersect on fixed batches -------------------------------------------------
    def trace_closest(self, o, d, tmax, which=0, kernel=0):
        n = o.shape[0]
        o, d, tmax = (np.ascontiguousarray(x, np.float32) for x in (o, d, tmax))
        out = dict(hit=np.zeros(n, np.uint8), t=np.zeros(n, np.float32), u=np.zeros(n, np.float32), v=np.zeros(n, np.float32), prim=np.zeros(n, np.uint32),
                   blas=np.zeros(n, np.uint32), inst=np.zeros(n, np.uint32), front=np.zeros(n, np.uint8), normal=np.zeros((n, 3), np.float32))
        self._ck(self.L.ptb_trace_closest(self.h, which, kernel, C.c_uint64(n), _p(o), _p(d), _p(tmax), _p(out["hit"]), _p(out["t"]), _p(out["u"]), _p(out["v"]),
                                          _p(out["prim"]), _p(out["blas"]), _p(out["inst"]), _p(out["front"]), _p(out["normal"])))
        return out

    def trace_closest_static(self, o, d, tmax, which=0):
        """ptb_trace_closest_static: the same batch trace without any device allocation (static scratch, <= STATIC_TRACE_MAX_RAYS rays)."""
        n = o.shape[0]
        o, d, tmax = (np.ascontiguousarray(x, np.float32) for x in (o, d, tmax))
        out = dict(hit=np.zeros(n, np.uint8), t=np.zeros(n, np.float32), u=np.zeros(n, np.float32), v=np.zeros(n, np.float32), prim=np.zeros(n, np.uint32),
                   blas=np.zeros(n, np.uint32), inst=np.zeros(n, np.uint32), front=np.zeros(n, np.uint8), normal=np.zeros((n, 3), np.float32))
        self._ck(self.L.ptb_trace_closest_static(self.h, which, C.c_uint64(n), _p(o), _p(d), _p(tmax), _p(out["hit"]), _p(out["t"]), _p(out["u"]), _p(out["v"]),
                                                 _p(out["prim"]), _p(out["blas"]), _p(out["inst"]), _p(out["front"]), _p(out["normal"])))
        return out

    def trace_any(self, o, d, tmax, which=0, kernel=0):
        n = o.shape[0]
        o, d, tmax = (np.ascontiguousarray(x, np.float32) for x in (o, d, tmax))
        hit = np.zeros(n, np.uint8)
        self._ck(self.L.ptb_trace_any(self.h, which, kernel, C.c_uint64(n), _p(o), _p(d), _p(tmax), _p(hit)))
        return dict(hit=hit)

    def trace_upload(self, o, d, tmax):
        o, d, tmax = (np.ascontiguousarray(x, np.float32) for x in (o, d, tmax))
        self._ck(self.L.ptb_trace_upload(self.h, C.c_uint64(o.shape[0]), _p(o), _p(d), _p(tmax)))
        self._ck(self.L.ptb_synchronize(self.h))
        self._tn = o.shape[0]

    def trace_resident(self, which=0, kernel=0, any_hit=False, repeats=1):
        ms = C.c_float()
        self._ck(self.L.ptb_trace_resident(self.h, which, kernel, int(any_hit), repeats, C.byref(ms)))
        return ms.value

    def trace_download(self):
        n = self._tn
        out = dict(hit=np.zeros(n, np.uint8), t=np.zeros(n, np.float32), prim=np.zeros(n, np.uint32), inst=np.zeros(n, np.uint32))
        self._ck(self.L.ptb_trace_download(self.h, _p(out["hit"]), _p(out["t"]), _p(out["prim"]), _p(out["inst"])))
        return out

    # ---- Camera / sampler probes -----------------------------------------------------------------------------
    def set_camera(self, origin, target, fov_deg, aspect):
        o = np.asarray(origin, np.float32)
        t = np.asarray(target, np.float32)
        self._ck(self.L.ptb_set_camera(self.h, _p(o), _p(t), C.c_float(fov_deg), C.c_float(aspect)))

    def camera_rays(self, st):
        st = np.ascontiguousarray(st, np.float32)
        n = st.shape[0]
        o = np.zeros((n, 3), np.float32)
        d = np.zeros((n, 3), np.float32)
        self._ck(self.L.ptb_camera_rays(self.h, C.c_uint64(n), _p(st), _p(o), _p(d)))
        return o, d

    def ss_sobol(self, index, seed):
        index = np.ascontiguousarray(index, np.uint32)
        seed = np.ascontiguousarray(seed, np.uint32)
        out = np.zeros((index.shape[0], 2), np.float32)
        self._ck(self.L.ptb_ss_sobol(self.h, C.c_uint64(index.shape[0]), _p(index), _p(seed), _p(out)))
        return out

    def rng_stream(self, seed, pixel, sample, n):
        out = np.zeros(n, np.uint32)
        self._ck(self.L.ptb_rng_stream(self.h, C.c_uint64(seed), pixel, sample, n, _p(out)))
        return out

    # ---- frame loop ------------------------------------------------------------------------------------------------
    def frame_begin(self, width, height, second_moment=False):
        self._ck(self.L.ptb_enable_second_moment(self.h, int(second_moment)))
        self._ck(self.L.ptb_frame_begin(self.h, width, height))
        self.width, self.height = width, height

    def render(self, sample_begin, sample_count, seed=1, max_bounces=1024, enable_nee=True):
        self._ck(self.L.ptb_render(self.h, sample_begin, sample_count, C.c_uint64(seed), max_bounces, int(enable_nee)))

    def render_async(self, sample_begin, sample_count, seed=1, max_bounces=1024, enable_nee=True):
        """ptb_render on a worker thread of the context; wait() joins it.  For two contexts on one device rendering alternate frames."""
        self._ck(self.L.ptb_render_async(self.h, sample_begin, sample_count, C.c_uint64(seed), max_bounces, int(enable_nee)))

    def wait(self):
        self._ck(self.L.ptb_wait(self.h))

    def read_accum(self, out=None):
        if out is None:
            out = np.zeros((self.height, self.width, 4), np.float32)
        self._ck(self.L.ptb_read_accum(self.h, _p(out)))
        return out

    def read_rgb8(self, flip_y=True):
        """Tonemapped 8-bit image of the accumulation buffer (write_image, image_helper.rs:37-58), top row first by default."""
        out = np.zeros((self.height, self.width, 3), np.uint8)
        self._ck(self.L.ptb_read_rgb8(self.h, _p(out), int(flip_y)))
        return out

    # ---- temporal accumulation: State::update (state.rs:505-579) ----------------------------------------------------
    def frame_next(self):
        """Next frame of the same size: clears the frame input, keeps the id history (main.rs:204-206)."""
        self._ck(self.L.ptb_frame_next(self.h))

    def view_projection(self):
        """(cam.matrix * cam.inv_projection).inverse() (main.rs:209-216) of the current camera, column-major [16]."""
        m = np.zeros(16, np.float32)
        self._ck(self.L.ptb_camera_view_projection(self.h, _p(m)))
        return m

    def state_reset(self, width, height):
        self._ck(self.L.ptb_state_reset(self.h, C.c_uint32(width), C.c_uint32(height)))

    def state_update(self, data, position, ids, inv_projection, last_inv_projection):
        """State::update with the reference's argument list (host arrays [H,W,4], [H,W,4], [H,W] u32, two Mat4)."""
        data = np.ascontiguousarray(data, np.float32)
        position = np.ascontiguousarray(position, np.float32)
        ids = np.ascontiguousarray(ids, np.uint32)
        H, W = data.shape[:2]
        if position.shape[:2] != (H, W) or ids.shape[:2] != (H, W):
            raise ValueError("data, position and id must have the same image size")
        a = np.ascontiguousarray(np.asarray(inv_projection, np.float32).reshape(16))
        b = np.ascontiguousarray(np.asarray(last_inv_projection, np.float32).reshape(16))
        self._ck(self.L.ptb_state_update(self.h, C.c_uint32(W), C.c_uint32(H), _p(data), _p(position), _p(ids), _p(a), _p(b)))
        self._state_shape = (H, W)

    def state_update_resident(self, inv_projection, last_inv_projection):
        """The same on the context's own frame buffers (accumulation buffer as frame input, aux position / id)."""
        a = np.ascontiguousarray(np.asarray(inv_projection, np.float32).reshape(16))
        b = np.ascontiguousarray(np.asarray(last_inv_projection, np.float32).reshape(16))
        self._ck(self.L.ptb_state_update_resident(self.h, _p(a), _p(b)))
        self._state_shape = (self.height, self.width)

    def state_read(self, velocity=False):
        H, W = self._state_shape
        out = np.zeros((H, W, 4), np.float32)
        vel = np.zeros((H, W, 2), np.float32) if velocity else None
        self._ck(self.L.ptb_state_read(self.h, _p(out), _p(vel)))
        return (out, vel) if velocity else out

    def state_timing(self, reset=True):
        """(total ms, launches) of the update kernels timed while set_profiling(True)"""
        ms, n = C.c_float(), C.c_uint32()
        self._ck(self.L.ptb_state_timing(self.h, C.byref(ms), C.byref(n), int(reset)))
        return ms.value, n.value

    def state_read_rgb8(self, flip_y=True):
        H, W = self._state_shape
        out = np.zeros((H, W, 3), np.uint8)
        self._ck(self.L.ptb_state_read_rgb8(self.h, _p(out), int(flip_y)))
        return out

    def read_second_moment(self):
        out = np.zeros((self.height, self.width, 4), np.float32)
        self._ck(self.L.ptb_read_second_moment(self.h, _p(out)))
        return out

    def read_aux(self, position=None, ids=None):
        if position is None:
            position = np.zeros((self.height, self.width, 4), np.float32)
        if ids is None:
            ids = np.zeros((self.height, self.width), np.uint32)
        self._ck(self.L.ptb_read_aux(self.h, _p(position), _p(ids)))
        return position, ids

    def accum_device_ptr(self):
        p = C.c_void_p()
        n = C.c_size_t()
        self._ck(self.L.ptb_accum_device_ptr(self.h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def bind_accum(self, ptr, nbytes):
        self._ck(self.L.ptb_bind_accum(self.h, C.c_void_p(ptr), C.c_size_t(nbytes)))

    def synchronize(self):
        self._ck(self.L.ptb_synchronize(self.h))

    def counters(self):
        out = np.zeros(8, np.uint64)
        self._ck(self.L.ptb_get_counters(self.h, _p(out)))
        d = {k: int(v) for k, v in zip(COUNTER_NAMES, out)}
        d["primary"] = d["paths"]  # every path starts with exactly one primary ray
        return d

    def reset_counters(self):
        self._ck(self.L.ptb_reset_counters(self.h))

    def set_profiling(self, on):
        self._ck(self.L.ptb_set_profiling(self.h, int(on)))

    def timing(self):
        t = np.zeros(8, np.float32)
        n = np.zeros(8, np.uint32)
        self._ck(self.L.ptb_get_timing(self.h, _p(t), _p(n)))
        names = ["render", "init", "extend", "shade", "nee", "tail", "reserved6", "reserved7"]
        return {k: (float(a), int(b)) for k, a, b in zip(names, t, n)}

    def iteration_log(self, capacity=8192):
        a = np.zeros(capacity, np.uint32)
        n = C.c_uint32()
        self._ck(self.L.ptb_get_iteration_log(self.h, capacity, _p(a), C.byref(n)))
        return a[: n.value]

    def launch_log(self, capacity=1 << 16):
        """(stage, microseconds) arrays of every launch of the last render that ran with profiling on."""
        st = np.zeros(capacity, np.int32)
        us = np.zeros(capacity, np.float32)
        n = C.c_uint32()
        self._ck(self.L.ptb_get_launch_log(self.h, capacity, _p(st), _p(us), C.byref(n)))
        return st[: n.value], us[: n.value]

    def set_option(self, name, value):
        self._ck(self.L.ptb_set_option(self.h, name.encode(), int(value)))


# ----------------------------------------------------------------------------------------------------------------------
# Mirror of the reference's own types (same names, argument meaning and error behaviour) above the C ABI
# ----------------------------------------------------------------------------------------------------------------------
class Lambertian:
    @staticmethod
    def new(albedo):  # material.rs:99
        return _scenes.material(_scenes.LAMBERTIAN, albedo)


class Emissive:
    @staticmethod
    def new(emitted):  # material.rs:126
        return _scenes.material(_scenes.EMISSIVE, emitted)


class Specular:
    @staticmethod
    def new(colour):  # material.rs:146
        return _scenes.material(_scenes.SPECULAR, colour)


class GGX:
    @staticmethod
    def new_metal(colour, roughness):  # material.rs:290
        return _scenes.material(_scenes.GGX_METAL, colour, roughness=roughness)

    @staticmethod
    def new_dielectric(colour, roughness, ior, volume=None):  # material.rs:305
        return _scenes.material(_scenes.GGX_DIELECTRIC, colour, roughness=roughness, ior=ior, volume=volume)


class Dielectric:
    @staticmethod
    def new(colour, ior, volume=None):  # material.rs:475
        return _scenes.material(_scenes.DIELECTRIC, colour, ior=ior, volume=volume)


class Volume:
    @staticmethod
    def new(absorption, k, c, g):  # volume.rs:136
        return (tuple(absorption), float(k), float(c), float(g))


class Model:
    """Model::new(file_path, material, matrices) (model.rs:36-51).  `mesh` is either a path to an OBJ file (loaded with the
    reference's rules, see obj.py) or a float32 [ntri,3,6] vertex array.  Asserts unit scale like the reference."""

    def __init__(self, mesh, material, matrices):
        if isinstance(mesh, (str, os.PathLike)):
            from .obj import load_obj
            mesh = load_obj(mesh)
        self.verts = np.ascontiguousarray(mesh, np.float32)
        self.material = material
        self.matrices = [tuple(float(x) for x in m) for m in matrices]
        for m in self.matrices:
            a = np.array(m[:9], np.float32).reshape(3, 3)
            scale = np.sqrt((a * a).sum(axis=1, dtype=np.float32)).astype(np.float32)
            assert np.all(scale == np.float32(1.0)), "Model matrix can only contain translation and rotation"


class Camera:
    """Camera::new(origin, target, fov, aspect_ratio, _, _) (camera.rs:17); aperture and focus distance are ignored upstream."""

    def __init__(self, origin, target, fov, aspect_ratio, _aperture=0.0, _focus=0.0):
        self.origin, self.target, self.fov, self.aspect_ratio = tuple(origin), tuple(target), float(fov), float(aspect_ratio)


class Scene:
    """Scene::new(models) (scene.rs:21-35) on one GPU; `render` is the pixel loop of main.rs:181-207 for `frames` frames."""

    def __init__(self, models, device=0):
        chunks, blas, inst, n = [], [], [], 0
        for i, m in enumerate(models):
            blas.append((n, int(m.verts.shape[0]), m.material))
            chunks.append(m.verts)
            n += int(m.verts.shape[0])
            inst += [(i, mm) for mm in m.matrices]
        self.sd = _scenes.SceneData("user", np.concatenate(chunks), blas, inst, (0, 0, 1), (0, 0, 0), 60.0, 16, 9, 1, 1024)
        self.ctx = Context(device)
        self.ctx.upload(self.sd)

    def render(self, camera: Camera, width, height, frames, seed=1, max_bounces=1024, enable_nee=True, first_frame=0):
        c = self.ctx
        c.set_camera(camera.origin, camera.target, camera.fov, camera.aspect_ratio)
        if (c.width, c.height) != (width, height) or first_frame == 0:
            c.frame_begin(width, height)
        c.render(first_frame, frames, seed=seed, max_bounces=max_bounces, enable_nee=enable_nee)
        data = c.read_accum()
        position, ids = c.read_aux()
        return data, position, ids
