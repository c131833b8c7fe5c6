"""Sample-partitioned rendering across the GPUs of one box: the binding of ptb_group_* (include/ptb.h, csrc/group.cu).

The reference is a single process on CPU threads; its frame loop (main.rs:181-207) is a pure map over (pixel, sample) and the
accumulation (accumulate.wgsl:23) a commutative float sum, so the path shards by *global sample index*: every GPU holds the
full scene, renders a contiguous block of a batch's samples for all pixels, and the W x H x 4 accumulation buffers are summed
onto the root with ONE ncclReduce per batch — issued by libptb.so itself (it links libnccl), on each member's render stream.
RNG streams are keyed by (seed, pixel, global sample index), so the union of the samples is the same for every group size.

Two ways to form a group, as in the C ABI:
  Group.create([0, 1, ...])          one process drives several devices (what a Rust host would do)
  Group.join(ctx, world, rank, id)   one process per GPU (torchrun): rank 0 calls unique_id(), the host ships the 128 bytes
                                     to the other ranks (torch.distributed broadcast in bench.py — plumbing only).
`sample_range` / `render_partitioned` / `reduce_accum` keep the partition arithmetic testable on CPU (gloo, world size 2):
sample_range must equal ptb_group_sample_range, which tests/test_multi_gloo.py checks against the library.
"""
from __future__ import annotations

import ctypes as C

import numpy as np


def sample_range(begin: int, count: int, rank: int, world: int):
    """Contiguous block of the global sample range [begin, begin+count) owned by `rank`: the first count % world ranks get one
    sample more.  Returns (first global sample index, number of samples); count may be 0 for ranks beyond the sample count.
    Same arithmetic as ptb_group_sample_range (csrc/group.cu)."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    base, extra = divmod(count, world)
    mine = base + (1 if rank < extra else 0)
    first = begin + rank * base + min(rank, extra)
    return first, mine


def lib_sample_range(begin: int, count: int, rank: int, world: int):
    """ptb_group_sample_range of libptb.so (host arithmetic; needs no device)."""
    from . import api
    first, mine = C.c_uint32(), C.c_uint32()
    rc = api.lib().ptb_group_sample_range(C.c_uint32(begin), C.c_uint32(count), int(rank), int(world), C.byref(first), C.byref(mine))
    if rc != api.PTB_OK:
        raise ValueError("rank out of range")
    return first.value, mine.value


def unique_id() -> bytes:
    """128-byte NCCL id for Group.join, made on one rank and shipped to the others by the host."""
    from . import api
    buf = (C.c_uint8 * 128)()
    L = api.lib()
    rc = L.ptb_group_unique_id(buf)
    if rc != api.PTB_OK:
        L.ptb_group_last_error.restype = C.c_char_p
        raise api.PtbError(rc, L.ptb_group_last_error(None).decode())
    return bytes(buf)


class Group:
    """A ptb_group: the members render disjoint sample blocks of every batch; reduce() sums them onto the root with NCCL."""

    def __init__(self, handle, contexts, owns):
        from . import api
        self.L = api.lib()
        self.L.ptb_group_last_error.restype = C.c_char_p
        self.L.ptb_group_last_error.argtypes = [C.c_void_p]
        self.L.ptb_group_destroy.restype = None
        self.L.ptb_group_destroy.argtypes = [C.c_void_p]
        self.h = handle
        self.contexts = contexts  # api.Context views of the local members
        self._owns = owns
        nl, nr = C.c_int(), C.c_int()
        self.L.ptb_group_size(self.h, C.byref(nl), C.byref(nr))
        self.n_local, self.n_ranks = nl.value, nr.value
        self.width = self.height = 0

    @classmethod
    def create(cls, devices):
        from . import api
        L = api.lib()
        L.ptb_group_last_error.restype = C.c_char_p
        L.ptb_group_last_error.argtypes = [C.c_void_p]
        L.ptb_group_context.restype = C.c_void_p
        L.ptb_group_context.argtypes = [C.c_void_p, C.c_int]
        devs = (C.c_int * len(devices))(*devices)
        h = C.c_void_p()
        rc = L.ptb_group_create(len(devices), devs, C.byref(h))
        if rc != api.PTB_OK:
            raise api.PtbError(rc, L.ptb_group_last_error(None).decode())
        ctxs = [api.Context.borrow(L.ptb_group_context(h, i), devices[i]) for i in range(len(devices))]
        return cls(h, ctxs, True)

    @classmethod
    def join(cls, ctx, world, rank, uid: bytes):
        from . import api
        L = api.lib()
        L.ptb_group_last_error.restype = C.c_char_p
        L.ptb_group_last_error.argtypes = [C.c_void_p]
        if len(uid) != 128:
            raise ValueError("unique id must be 128 bytes")
        buf = (C.c_uint8 * 128).from_buffer_copy(uid)
        h = C.c_void_p()
        rc = L.ptb_group_join(ctx.h, int(world), int(rank), buf, C.byref(h))
        if rc != api.PTB_OK:
            raise api.PtbError(rc, L.ptb_group_last_error(None).decode())
        return cls(h, [ctx], False)

    def close(self):
        if getattr(self, "h", None):
            self.L.ptb_group_destroy(self.h)
            self.h = None
            if self._owns:
                for c in self.contexts:
                    c.h = None  # destroyed with the group

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            from . import api
            raise api.PtbError(rc, self.L.ptb_group_last_error(self.h).decode())

    def upload(self, sd, build=True):
        from . import api
        verts, bd, inst = api.pack_scene(sd)
        self._ck(self.L.ptb_group_scene_upload(self.h, api._p(verts), C.c_uint64(verts.shape[0]), C.byref(bd), len(sd.blas), C.byref(inst), len(sd.instances)))
        for c in self.contexts:
            c.sd = sd
        if build:
            self._ck(self.L.ptb_group_build(self.h))

    def set_camera(self, origin, target, fov_deg, aspect):
        from . import api
        o, t = np.asarray(origin, np.float32), np.asarray(target, np.float32)
        self._ck(self.L.ptb_group_set_camera(self.h, api._p(o), api._p(t), C.c_float(fov_deg), C.c_float(aspect)))

    def set_option(self, name, value):
        self._ck(self.L.ptb_group_set_option(self.h, name.encode(), int(value)))

    def frame_begin(self, width, height):
        self._ck(self.L.ptb_group_frame_begin(self.h, width, height))
        self.width, self.height = width, height
        for c in self.contexts:
            c.width, c.height = width, height

    def render(self, sample_begin, sample_count, seed=1, max_bounces=1024, enable_nee=True):
        self._ck(self.L.ptb_group_render(self.h, sample_begin, sample_count, C.c_uint64(seed), max_bounces, int(enable_nee)))

    def reduce(self, root=0):
        self._ck(self.L.ptb_group_reduce(self.h, int(root)))

    def reduce_ms(self):
        ms = C.c_float()
        self._ck(self.L.ptb_group_reduce_ms(self.h, C.byref(ms)))
        return ms.value

    def read_accum(self, root=0, out=None):
        from . import api
        if out is None:
            out = np.zeros((self.height, self.width, 4), np.float32)
        self._ck(self.L.ptb_group_read_accum(self.h, int(root), api._p(out)))
        return out

    def read_aux(self, root=0, position=None, ids=None):
        from . import api
        if position is None:
            position = np.zeros((self.height, self.width, 4), np.float32)
        if ids is None:
            ids = np.zeros((self.height, self.width), np.uint32)
        self._ck(self.L.ptb_group_read_aux(self.h, int(root), api._p(position), api._p(ids)))
        return position, ids

    def counters(self):
        from . import api
        out = np.zeros(8, np.uint64)
        self._ck(self.L.ptb_group_get_counters(self.h, api._p(out)))
        return {k: int(v) for k, v in zip(api.COUNTER_NAMES, out)}

    def reset_counters(self):
        self._ck(self.L.ptb_group_reset_counters(self.h))


# ---- host-side mirror of the partition / reduce step on torch tensors (CPU tests with gloo; not the GPU product path) ----
def reduce_accum(acc, dst: int = 0, group=None):
    """Sum the accumulation buffers (torch tensor, W*H*4 float32: sum rgb, sample count) of all ranks onto `dst`, in place."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(acc, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return acc


def render_partitioned(render_fn, acc, begin: int, count: int, rank: int, world: int, dst: int = 0, group=None):
    """One frame batch on `world` ranks: `render_fn(first, n)` must add samples [first, first+n) of every pixel into `acc`.
    Afterwards `acc` on rank `dst` holds all `count` samples; on the other ranks it holds partial sums and must be cleared
    before the next batch.  This is what ptb_group_render + ptb_group_reduce do on the GPUs."""
    first, n = sample_range(begin, count, rank, world)
    if n:
        render_fn(first, n)
    return reduce_accum(acc, dst=dst, group=group)
