"""OBJ ingest with the reference's rules (load_obj, src/tlas/tlas_bvh/blas.rs:44-131) -> the [ntri, 3, 6] float32 vertex array
(position xyz, normal xyz) that ptb_scene_upload takes.  Host-side data-format code, as upstream.

Rules reproduced: `v`/`vn`/`f` only (everything else skipped, :122-126); 1-based indices with a dummy element 0 (:46-47);
`vn` normalised on load (:75); every face token must have three '/'-separated fields — `v//vn` or `v/vt/vn` (:90; upstream
panics on a bare `v`, here ValueError); a non-numeric-unsigned index is parsed as signed and added to the current array length
(:86-92); fan triangulation (refs[0], refs[i], refs[i+1]) (:97-103); normal index 0 -> the UN-normalised face normal
cross(p1-p0, p2-p0) of that fan triangle (:109-116).  f32 parsing is correctly rounded like Rust's `str::parse::<f32>`.
"""
from __future__ import annotations

from fractions import Fraction

import numpy as np

_F32 = np.float32


def parse_f32(tok: str) -> np.float32:
    """Correctly rounded decimal -> float32 (Rust's f32::from_str).  float() is correctly rounded to double; rounding that
    double to float32 is a second rounding, wrong only when the double sits exactly on a float32 tie while the decimal does
    not: resolved exactly with rationals."""
    d = float(tok)  # raises ValueError like .parse().unwrap() panics
    f = _F32(d)
    if not np.isfinite(f) or f == 0 or float(f) == d:
        return f
    lo, hi = (f, np.nextafter(f, _F32(np.inf))) if float(f) < d else (np.nextafter(f, _F32(-np.inf)), f)
    mid = (Fraction(float(lo)) + Fraction(float(hi))) / 2
    if Fraction(d) != mid:
        return f
    try:
        exact = Fraction(tok)
    except (ValueError, ZeroDivisionError):
        return f
    if exact == mid:
        return f  # a true tie: round-half-even is what the cast did
    return hi if exact > mid else lo


def _normalize(v):
    # glam Vec3A::normalize: v / sqrt((x*x + y*y) + z*z) in f32
    l2 = _F32(_F32(_F32(v[0] * v[0]) + _F32(v[1] * v[1])) + _F32(v[2] * v[2]))
    return (v / np.sqrt(l2, dtype=_F32)).astype(_F32)


def _cross(a, b):
    return np.array([_F32(a[1] * b[2]) - _F32(b[1] * a[2]), _F32(a[2] * b[0]) - _F32(b[2] * a[0]), _F32(a[0] * b[1]) - _F32(b[0] * a[1])], _F32)


def _index(field: str, length: int) -> int:
    try:
        v = int(field)
        if v >= 0 and field.strip()[0] != "-" and not field.startswith("+"):
            return v
    except ValueError:
        raise ValueError(f"bad OBJ index {field!r}") from None
    # parse::<usize>() failed (negative, or explicit '+' sign is accepted by Rust's usize parser — handled above): len + isize
    return length + int(field)


def load_obj_lines(lines) -> np.ndarray:
    positions = [np.zeros(3, _F32)]
    normals = [np.zeros(3, _F32)]
    out = []
    with np.errstate(all="ignore"):
        for line in lines:
            tokens = line.split()
            if not tokens:
                raise ValueError("empty line: the reference indexes tokens[0] and panics (blas.rs:58)")
            kw = tokens[0]
            if kw == "v":
                positions.append(np.array([parse_f32(t) for t in tokens[1:4]], _F32))
                if len(tokens) < 4:
                    raise ValueError("vertex needs three coordinates")
            elif kw == "vn":
                if len(tokens) < 4:
                    raise ValueError("normal needs three coordinates")
                normals.append(_normalize(np.array([parse_f32(t) for t in tokens[1:4]], _F32)))
            elif kw == "f":
                refs = []
                for tok in tokens[1:]:
                    fields = tok.split("/")
                    if len(fields) < 3:
                        raise ValueError(f"face token {tok!r} has no normal field: the reference requires v/vt/vn or v//vn (blas.rs:90)")
                    v = _index(fields[0], len(positions))
                    vn = _index(fields[2], len(normals))
                    if not (0 <= v < len(positions)) or not (0 <= vn < len(normals)):
                        raise ValueError(f"face token {tok!r}: index out of range")
                    refs.append((v, vn))
                if len(refs) < 2:
                    raise ValueError("face with fewer than two vertices (refs.len() - 1 underflows upstream)")
                for i in range(1, len(refs) - 1):
                    p0, p1, p2 = refs[0], refs[i], refs[i + 1]
                    for v, vn in (p0, p1, p2):
                        if vn != 0:
                            n = normals[vn]
                        else:
                            n = _cross(positions[p1[0]] - positions[p0[0]], positions[p2[0]] - positions[p0[0]])
                        out.append(np.concatenate([positions[v], n]))
            # comments, textures, groups, materials, unknown keywords: skipped
    if not out:
        return np.zeros((0, 3, 6), _F32)
    return np.asarray(out, _F32).reshape(-1, 3, 6)


def load_obj(path) -> np.ndarray:
    with open(path, "r") as f:
        return load_obj_lines(f.read().splitlines())
