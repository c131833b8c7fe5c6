"""ptb_trace_closest_static (csrc/noalloc.cu): the allocation-free variant of ptb_trace_closest — scratch in a file-scope static device
array instead of cudaMalloc'ed buffers — against the existing entry point, through the C ABI.  No counterpart in the reference.
Collected last on purpose; passes alone in a fresh process and
after the other tests."""
import numpy as np
import pytest

from conftest import bits

pytestmark = pytest.mark.gpu


def _same(a, b):
    for k in ("hit", "prim", "blas", "inst", "front"):
        assert np.array_equal(a[k], b[k]), k
    for k in ("t", "u", "v", "normal"):  # same kernel, same arithmetic: bit for bit (the tolerance of the other traversal tests on identical winners)
        assert np.array_equal(bits(a[k]), bits(b[k])), k


def test_static_scratch_trace_matches_allocating_entry_point(scenes):
    import path_tracer_b200 as ptb
    from path_tracer_b200 import api
    sd = scenes.scene_c3(grid=8)
    ctx = ptb.Context(0)
    ctx.upload(sd)
    # two calls in a row on different batches (different sizes: the second must not see the first one's leftovers), both TLASes
    o1, d1, t1 = scenes.incoherent_rays(sd, api.STATIC_TRACE_MAX_RAYS)
    o2, d2, t2 = scenes.incoherent_rays(sd, 12345, seed=77, finite_tmax=True)
    r1 = ctx.trace_closest_static(o1, d1, t1)
    r2 = ctx.trace_closest_static(o2, d2, t2)
    _same(r1, ctx.trace_closest(o1, d1, t1, kernel=0))
    _same(r2, ctx.trace_closest(o2, d2, t2, kernel=0))
    assert r1["hit"].sum() > 1000 and r2["hit"].sum() > 100
    _same(ctx.trace_closest_static(o2, d2, t2, which=1), ctx.trace_closest(o2, d2, t2, which=1, kernel=0))
    _same(ctx.trace_closest_static(o1, d1, t1), r1)  # repeatable: shares the array with the calls before it
    # a second context on the same device uses the same array (serialised here)
    ctx2 = ptb.Context(0)
    ctx2.upload(scenes.scene_c2())
    oc, dc, tc = scenes.incoherent_rays(ctx2.sd, 4097)
    _same(ctx2.trace_closest_static(oc, dc, tc), ctx2.trace_closest(oc, dc, tc, kernel=0))
    _same(ctx.trace_closest_static(o2, d2, t2), r2)


def test_static_scratch_limits(scenes):
    import path_tracer_b200 as ptb
    from path_tracer_b200 import api
    sd = scenes.scene_c2()
    ctx = ptb.Context(0)
    ctx.upload(sd)
    o, d, t = scenes.incoherent_rays(sd, api.STATIC_TRACE_MAX_RAYS + 1)
    with pytest.raises(ptb.PtbError) as e:
        ctx.trace_closest_static(o, d, t)
    assert e.value.code == api.PTB_ERR_INVALID and str(api.STATIC_TRACE_MAX_RAYS) in str(e.value)
    e0 = np.zeros((0, 3), np.float32)
    assert ctx.trace_closest_static(e0, e0, np.zeros(0, np.float32))["hit"].shape == (0,)
    fresh = ptb.Context(0)
    with pytest.raises(ptb.PtbError) as e:
        fresh.trace_closest_static(o[:8], d[:8], t[:8])
    assert e.value.code == api.PTB_ERR_NOT_BUILT
