"""BASELINE.json's full-size workload (diningroom stand-in, 1920x1080, depth 16) through size-independent properties (the
oracle / reference-kernel comparisons at this size live in tests/test_gpu_headline.py): the product walk against the reference-order walk on every
primary ray, invariance of the sums under the wave split and under sample sharding, device merge = single render,
and conservation of the sample count."""
import numpy as np
import pytest

from montecarlopathtracing_b200 import device, scenes, types as T

pytestmark = pytest.mark.gpu

W, H, D = 1920, 1080, 16


@pytest.fixture(scope="module")
def dining():
    return scenes.load("diningroom")


def _ctx(sc, **kw):
    ctx = device.Context(W, H, D, 1 << 20, seed_mode=1, run_seed=77, **kw)
    ctx.upload_scene(sc.triangles, sc.materials, sc.nodes)
    ctx.set_camera(sc.camera)
    return ctx


def test_primary_rays_fast_walk_equals_reference_walk(dining):
    with _ctx(dining) as ctx:
        ctx.generate_rays()
        ctx.intersect(device.TRAVERSAL_REFERENCE)
        ref, ref_closest = ctx.read_hits()
        ctx.intersect(device.TRAVERSAL_FAST)
        got, got_closest = ctx.read_hits()
    assert np.array_equal(ref_closest, got_closest)
    for f in ("normal", "point", "t", "material_id"):
        assert got[f].tobytes() == ref[f].tobytes(), f
    hit = ref_closest >= 0
    assert hit.mean() > 0.99
    # the centre column (d.x == 0 exactly) is the reference's 0/0 case: mostly misses, and the same misses
    col = hit.reshape(H, W)[:, W // 2]
    assert col.sum() < H // 4


def test_wave_split_and_sharding_do_not_change_the_sums(dining):
    with _ctx(dining, samples_in_flight=8) as a:
        a.render(8)
        one = a.read_sums()
        stats = a.stats()
    assert stats["paths"] == 8 * W * H
    assert 8 * W * H < stats["rays_traced"] <= 8 * W * H * D
    with _ctx(dining, samples_in_flight=3) as b:                  # waves of 3 + 3 + 2 samples
        b.render(8)
        split = b.read_sums()
    assert split.tobytes() == one.tobytes()
    with _ctx(dining, samples_in_flight=4) as r0, _ctx(dining, samples_in_flight=4) as r1:
        r0.render(5)
        r1.set_first_sample(5)
        r1.render(3)
        host_merge = r0.read_sums() + r1.read_sums()
        r0.merge_sums([r1])
        dev_merge = r0.read_sums()
    assert dev_merge.tobytes() == host_merge.tobytes()
    assert np.array_equal(dev_merge[6], one[6])                   # the counts of non-zero samples are integers: exact
    np.testing.assert_allclose(dev_merge, one, rtol=1e-12, atol=0)
    # conservation: every pixel counted at most 8 samples, and the image is not empty
    assert one[6].max() <= 8 and one[6].sum() > 0.5 * W * H
    assert np.isfinite(one).all() and (one[:3] >= 0).all()
