"""OBJ ingest: the binding of ptb_load_obj / ptb_load_obj_text (csrc/obj_loader.cpp), which implement the reference's load_obj
(src/tlas/tlas_bvh/blas.rs:44-131) as a product entry point -> the [ntri, 3, 6] float32 vertex array (position xyz, normal xyz)
that ptb_scene_upload takes.  Host-side data-format code, as upstream.

Rules (applied in C++, each cited there): `v`/`vn`/`f` only, everything else skipped (:122-126); 1-based indices with a dummy
element 0 (:46-47); `vn` normalised on load (:75); every face token must have three '/'-separated fields — `v//vn` or `v/vt/vn`
(:90; upstream panics on a bare `v`, here ValueError); an index that is not an unsigned decimal (a leading '+' is fine for Rust's
usize parser) is parsed as signed and added to the current array length (:86-92); fan triangulation (refs[0], refs[i], refs[i+1]),
so a face of one or two references yields no triangle (:97-103); normal index 0 -> the UN-normalised face normal
cross(p1-p0, p2-p0) of that fan triangle (:109-116).  f32 parsing is correctly rounded like Rust's `str::parse::<f32>`.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np


def _take(L, rc, verts, ntri):
    if rc != 0:
        L.ptb_load_obj_error.restype = C.c_char_p
        raise ValueError(L.ptb_load_obj_error().decode())
    n = ntri.value
    out = np.zeros((n, 3, 6), np.float32)
    if n:
        C.memmove(out.ctypes.data, verts, n * 72)
    L.ptb_free.argtypes = [C.c_void_p]
    L.ptb_free.restype = None
    L.ptb_free(verts)
    return out


def load_obj_text(text) -> np.ndarray:
    from . import api
    L = api.lib()
    b = text.encode() if isinstance(text, str) else bytes(text)
    verts, ntri = C.c_void_p(), C.c_uint64()
    rc = L.ptb_load_obj_text(b, C.c_size_t(len(b)), C.byref(verts), C.byref(ntri))
    return _take(L, rc, verts, ntri)


def load_obj_lines(lines) -> np.ndarray:
    """The lines of an OBJ file (without terminators), as BufRead::lines yields them upstream."""
    lines = list(lines)
    return load_obj_text("".join(ln + "\n" for ln in lines))


def load_obj(path) -> np.ndarray:
    from . import api
    L = api.lib()
    verts, ntri = C.c_void_p(), C.c_uint64()
    rc = L.ptb_load_obj(os.fspath(path).encode(), C.byref(verts), C.byref(ntri))
    return _take(L, rc, verts, ntri)
