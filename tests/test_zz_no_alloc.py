"""mcpt_trace_noalloc (include/mcpt.h): the allocation-free variant of mcpt_trace - device staging buffers and the walk's
work-fetch cursor in a static device array of the library instead of cudaMalloc'ed scratch - must return exactly what
the existing entry point returns (bit for bit, the tolerance of every intersection test here), on two calls in a row
(the array keeps the first call's contents), and must refuse batches beyond its stated maximum.  Collected last on
purpose; passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest

from montecarlopathtracing_b200 import device
from tests.util import adversarial_rays, hit_mismatches, random_rays

pytestmark = pytest.mark.gpu


def test_no_alloc_trace_equals_mcpt_trace_on_two_calls_in_a_row(cbox):
    with device.Context(64, 64, 8, 16) as ctx:
        ctx.upload_scene(cbox.triangles, cbox.materials, cbox.nodes)
        ctx.set_camera(cbox.camera)
        first = random_rays(device.TRACE_NOALLOC_MAX_RAYS, cbox.nodes)      # the full array
        second = adversarial_rays(cbox, 20_001)                              # fewer rays: stale records lie behind them
        second["state"][::7] = np.int32(-16777216)                           # 0xFF000000: terminated rays, skipped by intersect.cl:16
        for rays in (first, second):
            ref_hits, ref_closest = ctx.trace(rays, device.TRAVERSAL_FAST)   # the existing entry point
            live = (rays["state"].view(np.uint32) & 0xFF000000) == 0
            hits, closest = ctx.trace_noalloc(rays)
            assert np.array_equal(closest, ref_closest)
            assert hit_mismatches(hits[live], ref_hits[live]) == 0
            assert (closest[~live] == -1).all()
            assert not hits[~live].view(np.uint8).any()                      # cleared, not the previous call's records
            assert (closest[live] >= 0).mean() > 0.5
        # and again the first batch: same answer as before the smaller call
        hits, closest = ctx.trace_noalloc(first)
        ref_hits, ref_closest = ctx.trace(first, device.TRAVERSAL_FAST)
        assert np.array_equal(closest, ref_closest) and hit_mismatches(hits, ref_hits) == 0


def test_no_alloc_trace_rejects_batches_beyond_its_maximum(cbox):
    with device.Context(32, 32, 8, 16) as ctx:
        ctx.upload_scene(cbox.triangles, cbox.materials, cbox.nodes)
        rays = random_rays(device.TRACE_NOALLOC_MAX_RAYS + 1, cbox.nodes)
        with pytest.raises(device.McptError, match="exceed the maximum of 65536"):
            ctx.trace_noalloc(rays)
        hits, closest = ctx.trace_noalloc(rays[:0])                          # empty batch: nothing to do
        assert len(hits) == 0
