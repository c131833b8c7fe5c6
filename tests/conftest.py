import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "hostemu"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def ulp_diff(a, b):
    """distance in units of last place between two float32 arrays (same sign assumed for finite positives)"""
    ia = np.ascontiguousarray(a, np.float32).view(np.int32).astype(np.int64)
    ib = np.ascontiguousarray(b, np.float32).view(np.int32).astype(np.int64)
    return np.abs(ia - ib)


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def same_bits_or_nan(a, b):
    """float32 arrays equal bit for bit, except that any NaN matches any NaN (sign and payload of a NaN depend on operand order and
    on the machine: x86 returns the first NaN operand or the negative default NaN, the GPU the canonical 0x7fffffff)."""
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    return a.shape == b.shape and bool(np.all((bits(a) == bits(b)) | (np.isnan(a) & np.isnan(b))))


# SURVEY H2 expects ties within 1 ulp.  Measured (tests print the histogram, TIE_LOG collects it): most are, but on terrain-like meshes
# (c4) a few per 10^5 rays sit 2-4 ulps apart — two triangles sharing an edge, hit almost on that edge: the reference keeps whichever
# it tests LAST among candidates that pass `t in [t_min, t_max]` evaluated in the frame translated by the leaf's t_enter
# (primitive.rs:147-155), where t_max - t_estimate is rounded to the grid of the (larger) t_estimate; a candidate up to a few ulps
# NEARER than the current hit can fail that test, one a few ulps FARTHER can pass it.  Which of the two survives depends on visit
# order, so only the reference-order kernel reproduces it; for the 8-wide kernel such pairs are ties up to TIE_ULPS = 4 (<= 5e-7
# relative, the north star allows 1e-5 on t).  DESIGN.md section 2 lists examples.
TIE_ULPS = 4
TIE_LOG = []  # (ulps between the two candidates' t) of every classified tie of this session


def compare_closest(ref, got, oracle_scene, o, d, tmax, which=0, allow_ties=True):
    """Parity of a closest-hit batch against the oracle.

    Returns (n_exact, n_ties).  An index mismatch is a *tie* when both candidates, evaluated by the oracle exactly as a leaf
    visit would, are genuine hits whose t differ by at most TIE_ULPS ulps (<= 5e-7 relative; the north star allows 1e-5 on t).
    Such near-equal hits resolve by visit order upstream: the acceptance test of primitive.rs:122 runs in the pre-translated
    frame (t_max - t_estimate is rounded) and the leaf slab test clips against the running t_max, so a hit a few ulps nearer
    than the current t_max may be rejected.  The reference-order kernel reproduces even those; the 8-wide kernel visits in
    another order.  Anything else is an error."""
    assert np.array_equal(ref["hit"], got["hit"]), "hit/miss differs"
    h = ref["hit"] == 1
    mism = h & ((ref["prim"] != got["prim"]) | (ref["inst"] != got["inst"]))
    same = h & ~mism
    # identical winner -> bit-identical t, u, v, normal, front
    for k in ("t", "u", "v"):
        assert np.array_equal(bits(ref[k])[same], bits(got[k])[same]), f"{k} differs in bits on identical hits"
    assert np.array_equal(bits(ref["normal"])[same], bits(got["normal"])[same])
    assert np.array_equal(ref["front"][same], got["front"][same])
    assert np.array_equal(ref["blas"][same], got["blas"][same])
    idx = np.nonzero(mism)[0]
    if len(idx) and not allow_ties:
        raise AssertionError(f"{len(idx)} index mismatches where none are allowed")
    for i in idx:
        ok_r, t_r = oracle_scene.eval_candidate(ref["inst"][i], ref["prim"][i], o[i], d[i], tmax[i], which=which)
        ok_g, t_g = oracle_scene.eval_candidate(got["inst"][i], got["prim"][i], o[i], d[i], tmax[i], which=which)
        assert ok_r and ok_g, f"ray {i}: a reported winner is not a hit"
        assert ulp_diff(np.float32(t_r), np.float32(t_g)) <= TIE_ULPS, f"ray {i}: not a tie: {t_r} vs {t_g}"
        TIE_LOG.append(int(ulp_diff(np.float32(t_r), np.float32(t_g))))
        assert ulp_diff(ref["t"][i], got["t"][i]) <= TIE_ULPS
    return int(same.sum()), int(len(idx))


@pytest.fixture(scope="session")
def po():
    from oracle import po as _po
    _po.lib()
    return _po


@pytest.fixture(scope="session")
def scenes():
    from path_tracer_b200 import scenes as s
    return s
