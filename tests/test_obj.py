"""OBJ ingest (path_tracer_b200/obj.py) against the rules of load_obj, blas.rs:44-131 (expected values derived by hand from the
reference source: fan triangulation order, 1-based / negative indices, normalised vn, un-normalised fallback face normal)."""
import numpy as np
import pytest

from path_tracer_b200.obj import load_obj_lines, parse_f32

QUAD = """# a quad and a triangle
v 0 0 0
v 2 0 0
v 2 2 0
v 0 2 0
vn 0 0 3
vt 0.5 0.5
g grp
usemtl whatever
f 1//1 2//1 3//1 4//1
f -4/1/-1 -3/1/-1 -2/1/0
""".splitlines()


def test_fan_triangulation_and_indices():
    v = load_obj_lines(QUAD)
    assert v.shape == (3, 3, 6) and v.dtype == np.float32
    P = np.array([[0, 0, 0], [2, 0, 0], [2, 2, 0], [0, 2, 0]], np.float32)
    # refs[0], refs[i], refs[i+1]
    assert np.array_equal(v[0, :, :3], P[[0, 1, 2]]) and np.array_equal(v[1, :, :3], P[[0, 2, 3]])
    assert np.array_equal(v[:2, :, 3:], np.broadcast_to(np.float32([0, 0, 1]), (2, 3, 3)))  # vn normalised on load
    # negative indices count from the end (len includes the dummy element 0); normal index 0 -> raw cross(p1-p0, p2-p0)
    assert np.array_equal(v[2, :, :3], P[[0, 1, 2]])
    assert np.array_equal(v[2, 0, 3:], np.float32([0, 0, 1])) and np.array_equal(v[2, 1, 3:], np.float32([0, 0, 1]))
    assert np.array_equal(v[2, 2, 3:], np.float32([0, 0, 4]))  # un-normalised face normal


def test_reference_panics_become_errors():
    for bad in (["v 0 0 0", "v 1 0 0", "v 0 1 0", "f 1 2 3"], ["v 0 0 0", "", "v 1 1 1"], ["v 0 0"], ["v 0 0 0", "vn 0 0 1", "f 1//1 2//1 5//1"]):
        with pytest.raises(ValueError):
            load_obj_lines(bad)


def test_parse_f32_is_correctly_rounded():
    assert parse_f32("0.1") == np.float32(0.1) and parse_f32("-1e-3") == np.float32(-1e-3)
    # a decimal just above a float32 tie whose double rounds exactly onto the tie: double rounding would go down (even), correct is up
    lo = np.float32(1.0)
    hi = np.nextafter(lo, np.float32(2))
    s = "1.000000059604644775390625"  # exactly 1 + 2^-24, the tie between lo and hi
    assert float(s) == (float(lo) + float(hi)) / 2
    assert parse_f32(s + "0000000000000000000001") == hi   # float() lands on the tie, the cast alone would round to even (lo)
    assert parse_f32(s) == lo                               # a true tie rounds to even
    assert parse_f32("1.00000005960464477539062499999999") == lo


def test_model_loads_obj_through_the_interface(tmp_path):
    import path_tracer_b200 as ptb
    p = tmp_path / "quad.obj"
    p.write_text("\n".join(QUAD) + "\n")
    m = ptb.Model(str(p), ptb.Lambertian.new((0.5, 0.5, 0.5)), [(1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0)])
    assert m.verts.shape == (3, 3, 6)
    with pytest.raises(AssertionError):
        ptb.Model(str(p), ptb.Lambertian.new((0.5, 0.5, 0.5)), [(2, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0)])  # model.rs:42-43
