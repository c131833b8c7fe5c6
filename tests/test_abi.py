"""The C-ABI library loads, exports every symbol include/ptb.h declares, and refuses compute without a device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "ptb.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ptb_[a-z0-9_]+)\s*\(", src)))


def test_exports_match_header():
    import path_tracer_b200 as ptb
    from path_tracer_b200 import api
    L = ptb.lib()
    names = _declared()
    assert len(names) >= 35
    for n in names:
        assert hasattr(L, n), f"{n} declared in ptb.h but not exported by libptb.so"
    assert sorted(api.EXPORTS) == names
    assert L.ptb_abi_version() == 1


def test_struct_layout_matches_header():
    from path_tracer_b200 import api
    assert C.sizeof(api.Material) == 52 and C.sizeof(api.BlasDesc) == 60 and C.sizeof(api.Instance) == 52


def test_host_only_context_refuses_compute(scenes):
    import path_tracer_b200 as ptb
    ctx = ptb.Context(-1)
    sd = scenes.scene_c2()
    ctx.upload(sd)
    o, d, tm = scenes.incoherent_rays(sd, 8)
    for call in (lambda: ctx.trace_closest(o, d, tm), lambda: ctx.trace_any(o, d, tm), lambda: ctx.frame_begin(8, 8), lambda: ctx.camera_rays(np.zeros((1, 2), np.float32)),
                 lambda: ctx.rng_stream(1, 2, 3, 4)):
        with pytest.raises(ptb.PtbError) as e:
            call()
        assert e.value.code == ptb.api.PTB_ERR_NO_DEVICE


def test_validation_errors_mirror_reference_panics(scenes):
    import path_tracer_b200 as ptb
    ctx = ptb.Context(-1)
    sd = scenes.scene_c2()
    # no emissive model: Scene::new panics upstream (scene.rs:26 -> tlas_bvh.rs:58)
    import copy
    bad = copy.deepcopy(sd)
    bad.blas = [(f, c, dict(m, kind=scenes.LAMBERTIAN)) for f, c, m in bad.blas]
    with pytest.raises(ptb.PtbError) as e:
        ctx.upload(bad)
    assert e.value.code == ptb.api.PTB_ERR_INVALID and "emissive" in str(e.value)
    # non-unit scale: Model::new asserts (model.rs:43)
    bad = copy.deepcopy(sd)
    b0, m0 = bad.instances[1]
    bad.instances[1] = (b0, (2.0,) + tuple(m0[1:]))
    with pytest.raises(ptb.PtbError) as e:
        ctx.upload(bad)
    assert "translation and rotation" in str(e.value)
    # empty BLAS / out-of-order instances / not built
    bad = copy.deepcopy(sd)
    bad.blas[2] = (bad.blas[2][0], 0, bad.blas[2][2])
    with pytest.raises(ptb.PtbError):
        ctx.upload(bad)
    bad = copy.deepcopy(sd)
    bad.instances = bad.instances[::-1]
    with pytest.raises(ptb.PtbError):
        ctx.upload(bad)
    fresh = ptb.Context(-1)
    with pytest.raises(ptb.PtbError) as e:
        fresh.bvh_stats()
    assert e.value.code == ptb.api.PTB_ERR_NOT_BUILT
    # the Python mirror asserts like Model::new
    with pytest.raises(AssertionError):
        ptb.Model(sd.verts[:2], ptb.Lambertian.new((1, 1, 1)), [(2.0, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0)])


def test_static_scratch_entry_point_refuses_without_device(scenes):
    import path_tracer_b200 as ptb
    ctx = ptb.Context(-1)
    sd = scenes.scene_c2()
    ctx.upload(sd)
    o, d, tm = scenes.incoherent_rays(sd, 8)
    with pytest.raises(ptb.PtbError) as e:
        ctx.trace_closest_static(o, d, tm)
    assert e.value.code == ptb.api.PTB_ERR_NO_DEVICE
