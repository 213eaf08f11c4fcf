"""`yafgpu_trace_static` — the batch trace for callers that may not allocate device memory — against `yafgpu_trace`.

Collected last on purpose.  The entry point keeps its device scratch (ray copy, hit records,
work counter) in a file-scope `static __device__` array of libyafgpu.so instead of the memory pool, so two calls in a row reuse
the same bytes: the second call below uses other rays, fewer of them, and the other trace mode, and must not see the first
call's leftovers.  Results are compared bit for bit (the tolerance of every hit test in this suite: ids, t, u, v identical)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import parity_cases as pc  # noqa: E402

pytestmark = pytest.mark.gpu


def _same(a, b):
    return np.array_equal(a.view(np.uint8), b.view(np.uint8))


def test_static_scratch_trace_equals_pool_trace_on_two_calls_in_a_row():
    from libyafaray_b200 import Context, capi
    c = Context(0)
    try:
        y = pc.grid_mesh_scene(96, seed=11)
        desc, keep = y.scene_desc()
        c.upload(desc, keep)
        n_max = capi.TRACE_STATIC_MAX_RAYS
        r1 = pc.trace_rays(n_max, (-5, -5, -1), (5, 5, 2), seed=21)            # the maximum batch
        r2 = pc.trace_rays(20000, (-4, -4, 0.7), (4, 4, 3), seed=22)           # a shorter one: stale rays/hits beyond it must not matter
        r2[:, 6] = -np.abs(r2[:, 6]) - 0.1
        want1, want2 = c.trace(r1, capi.TRACE_CLOSEST), c.trace(r2, capi.TRACE_CLOSEST)
        want2s = c.trace(r2, capi.TRACE_SHADOW)
        got1 = c.trace_static(r1, capi.TRACE_CLOSEST)
        got2 = c.trace_static(r2, capi.TRACE_CLOSEST)
        got2s = c.trace_static(r2, capi.TRACE_SHADOW)
        got1b = c.trace_static(r1, capi.TRACE_CLOSEST)                          # and back to the large batch
        assert (want1["tri"] >= 0).sum() > 1000 and (want2["tri"] >= 0).sum() > 1000 and (want2s["tri"] >= 0).sum() > 100
        assert _same(got1, want1)
        assert _same(got2, want2)
        assert np.array_equal(got2s["tri"] >= 0, want2s["tri"] >= 0)           # any-hit: which blocker is found depends on the schedule
        assert _same(got1b, want1)
        # a second context on the same GPU shares the array (same process: serialised by the library)
        c2 = Context(0)
        try:
            c2.upload(desc, keep)
            assert _same(c2.trace_static(r2, capi.TRACE_CLOSEST), want2)
            assert _same(c.trace_static(r1, capi.TRACE_CLOSEST), want1)
        finally:
            c2.close()
    finally:
        c.close()


def test_static_scratch_trace_rejects_what_does_not_fit():
    from libyafaray_b200 import Context, capi
    c = Context(0)
    try:
        y = pc.grid_mesh_scene(16, seed=3)
        desc, keep = y.scene_desc()
        rays = pc.trace_rays(capi.TRACE_STATIC_MAX_RAYS + 1, (-5, -5, -1), (5, 5, 2), seed=23)
        with pytest.raises(capi.YafGpuError, match="no scene uploaded"):
            c.trace_static(rays[:16])
        c.upload(desc, keep)
        with pytest.raises(capi.YafGpuError, match="exceed the static scratch capacity of 65536"):
            c.trace_static(rays)
        with pytest.raises(capi.YafGpuError, match="mode must be"):
            c.trace_static(rays[:16], capi.TRACE_SHADOW_TRANSP)
        assert len(c.trace_static(rays[:0])) == 0                               # empty batch: nothing to do, no error
        h = c.trace_static(rays[:-1])                                           # exactly the maximum
        assert np.array_equal(h.view(np.uint8), c.trace(rays[:-1]).view(np.uint8))
    finally:
        c.close()
