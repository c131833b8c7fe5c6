"""OBJ ingest (ptb_load_obj / ptb_load_obj_text, csrc/obj_loader.cpp, through path_tracer_b200/obj.py) against the rules of load_obj,
blas.rs:44-131: values derived by hand from the reference source (fan triangulation order, 1-based / negative indices, normalised
vn, un-normalised fallback face normal) and, on generated OBJ text, against the oracle's literal restatement (oracle/po_obj.hpp)."""
from fractions import Fraction

import numpy as np
import pytest

from conftest import bits
from path_tracer_b200.obj import load_obj_lines, load_obj_text


def parse_f32(tok):
    """One OBJ coordinate through the product parser (a `v` line and a degenerate face that echoes it back)."""
    v = load_obj_text(f"v {tok} 0 0\nv 0 0 0\nv 0 0 0\nvn 0 0 1\nf 1//1 2//1 3//1\n")
    return v[0, 0, 0]

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


def test_reference_panics_become_errors(po):
    for bad in (["v 0 0 0", "v 1 0 0", "v 0 1 0", "f 1 2 3"], ["v 0 0 0", "", "v 1 1 1"], ["v 0 0"], ["v 0 0 0", "vn 0 0 1", "f 1//1 2//1 5//1"],
                ["v 0 0 0", "f"], ["v 0x10 0 0"], ["v 1e 0 0"], ["v . 0 0"], ["v 0 0 0", "vn 0 0 1", "f 1//1 1//1 a//1"], ["v 0 0 0", "vn 0 0 1", "f 1//1 1//1 -9//1"]):
        with pytest.raises(ValueError):
            load_obj_lines(bad)
        with pytest.raises(ValueError):
            po.load_obj_text("".join(x + "\n" for x in bad))


def test_short_faces_and_signs_follow_the_reference(po):
    """A face of one or two references is an empty fan (1..(refs.len()-1), blas.rs:97) — no triangle, no error, even when its
    indices point nowhere (they are never dereferenced); `+3` is a valid usize for Rust's parser; `-0` is not (-> len + 0, out of
    range only if used)."""
    base = ["v 0 0 0", "v 1 0 0", "v 0 1 0", "vn 0 0 1"]
    for lines in (base + ["f 1//1"], base + ["f 1//1 2//1"], base + ["f 99//99 98//98"]):
        assert load_obj_lines(lines).shape == (0, 3, 6)
        assert po.load_obj_text("".join(x + "\n" for x in lines)).shape == (0, 3, 6)
    a = load_obj_lines(base + ["f +1//+1 +2//1 3//1"])
    b = load_obj_lines(base + ["f 1//1 2//1 3//1"])
    assert np.array_equal(a, b) and a.shape == (1, 3, 6)
    with pytest.raises(ValueError):
        load_obj_lines(base + ["f -0//1 2//1 3//1"])


def test_generated_obj_against_the_oracle(po):
    """Random OBJ text — positive / negative / '+' indices, v//vn and v/vt/vn, missing normals (index 0), polygons of 3..7 vertices,
    comments, vt / g / usemtl noise, CRLF and tab separators, numbers in every accepted spelling — product against the oracle's
    statement-by-statement restatement, bit for bit."""
    rng = np.random.default_rng(7)
    spell = [lambda x: repr(float(x)), lambda x: "%.9g" % x, lambda x: "%e" % x, lambda x: ("+" if x >= 0 else "") + "%.7f" % x, lambda x: "%d." % int(x), lambda x: ("%.4f" % (x % 1)).lstrip("0") or "0"]
    for trial in range(6):
        lines, nv, nn = ["# generated"], 0, 0
        for _ in range(int(rng.integers(60, 200))):
            r = rng.random()
            if r < 0.45 or nv < 3:
                c = rng.normal(0, 50, 3)
                lines.append("v " + " ".join(spell[int(rng.integers(len(spell)))](x) for x in c))
                nv += 1
            elif r < 0.6:
                c = rng.normal(0, 1, 3) + 0.1
                lines.append("vn\t" + "\t".join(spell[int(rng.integers(len(spell) - 2))](x) for x in c))
                nn += 1
            elif r < 0.7:
                lines.append(["vt 0.5 0.25", "g group", "usemtl m", "# note", "o thing", "s off"][int(rng.integers(6))])
            else:
                k = int(rng.integers(3, 8))
                toks = []
                for _ in range(k):
                    vi = int(rng.integers(1, nv + 1))
                    vs = [str(vi), str(vi - nv - 1), "+" + str(vi)][int(rng.integers(3))]
                    if nn and rng.random() < 0.7:
                        ni = int(rng.integers(1, nn + 1))
                        ns = [str(ni), str(ni - nn - 1)][int(rng.integers(2))]
                    else:
                        ns = "0"
                    toks.append(vs + ["//", "/7/", "/x/"][int(rng.integers(3))] + ns)
                lines.append("f " + " ".join(toks))
        text = ("\r\n" if trial % 2 else "\n").join(lines) + "\n"
        got, ref = load_obj_text(text), po.load_obj_text(text)
        assert got.shape == ref.shape and got.shape[0] > 10
        assert np.array_equal(bits(got), bits(ref)), f"trial {trial}"


def test_obj_file_entry_point(tmp_path, po):
    from path_tracer_b200.obj import load_obj
    p = tmp_path / "quad.obj"
    p.write_text("\n".join(QUAD) + "\n")
    assert np.array_equal(bits(load_obj(p)), bits(po.load_obj_text(p.read_text())))
    with pytest.raises(ValueError):
        load_obj(tmp_path / "missing.obj")


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
