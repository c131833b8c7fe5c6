"""ctypes wrapper of tests/hostemu/libhostemu.so — TEST-ONLY host build of the product's traversal code."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        subprocess.check_call(["make", "-C", _HERE, "-s"], stderr=subprocess.DEVNULL)
        _LIB = C.CDLL(os.path.join(_HERE, "libhostemu.so"))
        _LIB.emu_create.restype = C.c_void_p
    return _LIB


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class EmuScene:
    def __init__(self, sd, pack_scene, threads=0):
        verts, bd, inst = pack_scene(sd)
        self._keep = (verts, bd, inst)
        err = C.create_string_buffer(512)
        self.h = lib().emu_create(_p(verts), C.c_uint64(verts.shape[0]), C.byref(bd), len(sd.blas), C.byref(inst), len(sd.instances), threads, err, 512)
        if not self.h:
            raise ValueError(err.value.decode())

    def close(self):
        if self.h:
            lib().emu_destroy(C.c_void_p(self.h))
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def stats(self):
        out = np.zeros(8, np.uint64)
        lib().emu_stats(C.c_void_p(self.h), _p(out))
        return out

    def trace(self, o, d, tmax, which=0, kernel=0, any_hit=False, threads=0):
        n = o.shape[0]
        o, d, tmax = (np.ascontiguousarray(x, np.float32) for x in (o, d, tmax))
        out = dict(hit=np.zeros(n, np.uint8), t=np.zeros(n, np.float32), u=np.zeros(n, np.float32), v=np.zeros(n, np.float32), prim=np.zeros(n, np.uint32),
                   blas=np.zeros(n, np.uint32), inst=np.zeros(n, np.uint32), front=np.zeros(n, np.uint8), normal=np.zeros((n, 3), np.float32))
        lib().emu_trace(C.c_void_p(self.h), which, kernel, int(any_hit), C.c_uint64(n), _p(o), _p(d), _p(tmax), _p(out["hit"]), _p(out["t"]), _p(out["u"]), _p(out["v"]),
                        _p(out["prim"]), _p(out["blas"]), _p(out["inst"]), _p(out["front"]), _p(out["normal"]), threads)
        return out


def trace_mr(es, o, d, tmax, which=0, any_hit=False, K=3, free_quorum=24, qn=16, ql=12, qi=12):
    """multiray.cuh (K rays per lane, gated stages) on one emulated warp; returns hit/t/u/v/prim/inst and the stage statistics"""
    n = o.shape[0]
    o, d, tmax = (np.ascontiguousarray(x, np.float32) for x in (o, d, tmax))
    out = dict(hit=np.zeros(n, np.uint8), t=np.zeros(n, np.float32), u=np.zeros(n, np.float32), v=np.zeros(n, np.float32), prim=np.zeros(n, np.uint32),
               inst=np.zeros(n, np.uint32), stats=np.zeros(8, np.uint64))
    lib().emu_trace_mr(C.c_void_p(es.h), which, int(any_hit), C.c_uint64(n), _p(o), _p(d), _p(tmax), K, free_quorum, qn, ql, qi, _p(out["hit"]), _p(out["t"]), _p(out["u"]),
                       _p(out["v"]), _p(out["prim"]), _p(out["inst"]), _p(out["stats"]))
    return out


TILE_LIST_STRIDE = 16


def tile_lists(es, sd, W, H):
    """tile_list_build (wavefront.h) for every 8 x 4 pixel tile of a W x H frame: (lists[n_tiles, 16, 2] uint32, slot -> pixel map,
    instance -> TLAS leaf-header index)"""
    m, eye = camera(sd.cam_origin, sd.cam_target, sd.fov_deg, np.float32(W) / np.float32(H))
    n_tiles = (W // 8) * (H // 4)
    lists = np.zeros((n_tiles, TILE_LIST_STRIDE, 2), np.uint32)
    s2p = np.zeros(W * H, np.uint32)
    lib().emu_tile_lists(C.c_void_p(es.h), _p(m), _p(eye), W, H, _p(lists), _p(s2p))
    n_inst = len(sd.instances)
    leaf = np.zeros(n_inst, np.uint32)
    lib().emu_instance_leaf_headers(C.c_void_p(es.h), _p(leaf), n_inst)
    return lists, s2p, leaf


def trace_list(es, lists, tile_of, o, d):
    """closest hit of camera rays from their tile's list (wide_begin_list, traverse.h), ray i in tile tile_of[i]"""
    n = o.shape[0]
    o, d = (np.ascontiguousarray(x, np.float32) for x in (o, d))
    tile_of = np.ascontiguousarray(tile_of, np.uint32)
    out = dict(hit=np.zeros(n, np.uint8), t=np.zeros(n, np.float32), u=np.zeros(n, np.float32), v=np.zeros(n, np.float32), prim=np.zeros(n, np.uint32), inst=np.zeros(n, np.uint32))
    lib().emu_trace_list(C.c_void_p(es.h), _p(lists), C.c_uint64(n), _p(tile_of), _p(o), _p(d), _p(out["hit"]), _p(out["t"]), _p(out["u"]), _p(out["v"]), _p(out["prim"]), _p(out["inst"]))
    return out


def camera(origin, target, fov_deg, aspect):
    m = np.zeros(16, np.float32)
    eye = np.zeros(3, np.float32)
    o = np.asarray(origin, np.float32)
    t = np.asarray(target, np.float32)
    lib().emu_camera(_p(o), _p(t), C.c_float(fov_deg), C.c_float(aspect), _p(m), _p(eye))
    return m, eye


def camera_rays(origin, target, fov_deg, aspect, st):
    m, eye = camera(origin, target, fov_deg, aspect)
    st = np.ascontiguousarray(st, np.float32)
    n = st.shape[0]
    o = np.zeros((n, 3), np.float32)
    d = np.zeros((n, 3), np.float32)
    lib().emu_camera_rays(_p(m), _p(eye), C.c_uint64(n), _p(st), _p(o), _p(d))
    return o, d


def ss_sobol(index, seed):
    index = np.ascontiguousarray(index, np.uint32)
    seed = np.ascontiguousarray(seed, np.uint32)
    out = np.zeros((index.shape[0], 2), np.float32)
    lib().emu_ss_sobol(C.c_uint64(index.shape[0]), _p(index), _p(seed), _p(out))
    return out


def rng_stream(seed, pixel, sample, n):
    out = np.zeros(n, np.uint32)
    lib().emu_rng_stream(C.c_uint64(seed), pixel, sample, n, _p(out))
    return out


def render(es, sd, width, height, sample_begin=0, sample_count=1, seed=1, max_bounces=None, enable_nee=True, extend_kernel=0):
    m, eye = camera(sd.cam_origin, sd.cam_target, sd.fov_deg, np.float32(width) / np.float32(height))
    accum = np.zeros((height, width, 4), np.float32)
    pos = np.zeros((height, width, 4), np.float32)
    ids = np.zeros((height, width), np.uint32)
    cnt = np.zeros(5, np.uint64)
    mb = sd.max_bounces if max_bounces is None else max_bounces
    lib().emu_render(C.c_void_p(es.h), _p(m), _p(eye), width, height, sample_begin, sample_count, C.c_uint64(seed), mb, int(enable_nee), extend_kernel, _p(accum), _p(pos), _p(ids), _p(cnt))
    return dict(accum=accum, position=pos, id=ids, counters=dict(extend=int(cnt[0]), shadow=int(cnt[1]), probe=int(cnt[2]), paths=int(cnt[3]), iters=int(cnt[4])))


def state_update(data, position, ids, inv_projection, last_inv_projection, history):
    """temporal.cuh bodies on the host; returns (new accumulation, velocity or None)"""
    data = np.ascontiguousarray(data, np.float32)
    position = np.ascontiguousarray(position, np.float32)
    ids = np.ascontiguousarray(ids, np.uint32)
    history = np.ascontiguousarray(history, np.float32)
    H, W = data.shape[:2]
    a = np.asarray(inv_projection, np.float32).reshape(16)
    b = np.ascontiguousarray(np.asarray(last_inv_projection, np.float32).reshape(16))
    moved = not bool(np.all(a == b))
    out = np.zeros((H, W, 4), np.float32)
    vel = np.zeros((H, W, 2), np.float32)
    lib().emu_state_update(W, H, _p(data), _p(position), _p(ids), _p(b), int(moved), _p(history), _p(out), _p(vel))
    return out, (vel if moved else None)
