"""Edge cases of the hot path on the GPU: the orthographic generator, grazing-incidence rays (the pruning margin's
theoretical gap), checkpoint / resume of the progressive state, the tagged scene upload and the
seed_mode 0 sharding guard."""
import numpy as np
import pytest

from montecarlopathtracing_b200 import device, types as T
from oracle import oracle_api as O
from tests.util import hit_mismatches

pytestmark = pytest.mark.gpu


def _ctx(sc, w, h, **kw):
    kw.setdefault("max_depth", 8)
    kw.setdefault("max_attempt", 1 << 20)
    ctx = device.Context(w, h, **kw)
    ctx.upload_scene(sc.triangles, sc.materials, sc.nodes)
    ctx.set_camera(sc.camera)
    return ctx


def test_orthographic_camera_bit_exact(cbox):
    """rayGenerator.cl:23-28 (cameraType != 0): origins spread over the image plane, one shared direction."""
    w, h = 160, 96
    cam = cbox.camera.copy()
    cam["camera_type"] = 1
    cam["arg"] = 400.0                       # the orthographic branch scales the offsets by `arg`
    ref = O.generate_rays(cam, w, h)
    if O.refcl() is not None:
        assert O.refcl_generate_rays(cam, w, h).tobytes() == ref.tobytes()
    ref_hits, ref_closest = O.intersect(ref, cbox.nodes, cbox.triangles)
    with device.Context(w, h, 8, 16) as ctx:
        ctx.upload_scene(cbox.triangles, cbox.materials, cbox.nodes)
        ctx.set_camera(cam)
        ctx.generate_rays()
        assert ctx.read_rays().tobytes() == ref.tobytes()
        ctx.intersect(device.TRAVERSAL_FAST)
        hits, closest = ctx.read_hits()
        assert np.array_equal(closest, ref_closest) and hit_mismatches(hits, ref_hits) == 0
        assert (closest >= 0).mean() > 0.3
        # and through the fused render (the wave generator has its own call site of generate_ray)
        ctx2 = device.Context(w, h, 6, 16, seed_mode=0, run_seed=3)
        ctx2.upload_scene(cbox.triangles, cbox.materials, cbox.nodes)
        ctx2.set_camera(cam)
        ctx2.render(2)
        g_hist, g_counts = ctx2.read_history()
        ctx2.close()
    r = O.render(cam, w, h, cbox.nodes, cbox.triangles, cbox.materials, 6, 16, 2, seed_mode=0, run_seed=3)
    assert g_hist.tobytes() == r["hist"].tobytes() and np.array_equal(g_counts, r["counts"])


def grazing_rays(sc, n, seed=5, max_cos=1e-4):
    """Rays that meet a triangle's plane at |cos(theta)| <= max_cos: the direction lies (almost) in the plane, aimed
    at the centroid from a point of the plane's own neighbourhood.  This is where the triangle solve's t can undershoot
    the leaf box's entry distance - the assumption behind the pruning margin (mcpt_traverse.cuh, PRUNE_REL)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    tris = sc.triangles
    t = rng.integers(0, len(tris), n)
    v = tris["v"][t][:, :, :3].astype(np.float64)
    nrm = tris["normal"][t].astype(np.float64)
    cen = v.mean(axis=1)
    e = v[:, 1] - v[:, 0]
    e /= np.maximum(np.linalg.norm(e, axis=1, keepdims=True), 1e-30)
    f = np.cross(nrm, e)
    ang = 2 * np.pi * rng.random(n)
    inplane = np.cos(ang)[:, None] * e + np.sin(ang)[:, None] * f          # unit vector in the triangle's plane
    c = (rng.random(n) * 2 - 1) * max_cos                                  # signed cosine against the normal
    d = inplane * np.sqrt(1 - c * c)[:, None] + nrm * c[:, None]
    dist = np.exp(rng.uniform(np.log(0.05), np.log(300.0), n))             # from grazing-near to across the room
    org = cen - d * dist[:, None]
    rays = np.zeros(n, T.RAY)
    rays["origin"] = org.astype(np.float32)
    rays["direction"] = d.astype(np.float32)
    rays["pixel"] = np.arange(n, dtype=np.int32)
    return rays


@pytest.mark.parametrize("name", ["cbox", "mis"])
def test_grazing_incidence_rays_bit_exact(name, request):
    sc = request.getfixturevalue(name)
    for max_cos in (1e-4, 1e-6, 0.0):
        rays = grazing_rays(sc, 40_000, max_cos=max_cos)
        ref_hits, ref_closest = O.intersect(rays, sc.nodes, sc.triangles)
        with _ctx(sc, 64, 64) as ctx:
            for trav in (device.TRAVERSAL_FAST, device.TRAVERSAL_CANONICAL):
                hits, closest = ctx.trace(rays, trav)
                assert np.array_equal(closest, ref_closest), (max_cos, trav, int((closest != ref_closest).sum()))
                assert hit_mismatches(hits, ref_hits) == 0, (max_cos, trav)
        assert (ref_closest >= 0).mean() > 0.5


@pytest.mark.parametrize("mode", [0, 1])
def test_checkpoint_and_resume_continue_bit_for_bit(cbox, mode):
    """colorout.cpp:23-24,49-50,63-68 keeps hist / counts / seeds on the device for the whole run; here they can be
    written to disk and a fresh context continues exactly where the first one stopped."""
    w, h = 96, 64
    kw = dict(max_depth=8, max_attempt=1 << 20, seed_mode=mode, run_seed=17, samples_in_flight=3 if mode else 0)
    with _ctx(cbox, w, h, **kw) as a:
        a.render(11)
        full_hist, full_counts = a.read_history()
        full_sums = a.read_sums()
        full_seeds = a.read_seeds()
    with _ctx(cbox, w, h, **kw) as b:
        b.render(4)
        ck = b.checkpoint()
    with _ctx(cbox, w, h, **kw) as c:
        c.restore(ck, mode)
        c.render(7)
        hist, counts = c.read_history()
        assert hist.tobytes() == full_hist.tobytes() and np.array_equal(counts, full_counts)
        assert c.read_sums().tobytes() == full_sums.tobytes()
        if mode == 0:
            assert np.array_equal(c.read_seeds(), full_seeds)
        else:
            assert c.first_sample() == 11


def test_tagged_upload_skips_the_comparison_but_not_the_copy(cbox):
    w = h = 64
    with _ctx(cbox, w, h, seed_mode=0, run_seed=2) as ctx:
        ctx.render(2)
        a, _ = ctx.read_history()
        mats = cbox.materials.copy()
        mats["kd"][:, :3] *= 0.5                   # same geometry, other materials: the arrays are copied again
        ctx.upload_scene(cbox.triangles, mats, cbox.nodes, geometry_tag=7)
        ctx.upload_scene(cbox.triangles, mats, cbox.nodes, geometry_tag=7)
        ctx.reset_accumulation()
        ctx.render(2)
        b, _ = ctx.read_history()
    ref = O.render(cbox.camera, w, h, cbox.nodes, cbox.triangles, mats, 8, 1 << 20, 2, seed_mode=0, run_seed=2)
    assert b.tobytes() == ref["hist"].tobytes()
    assert a.tobytes() != b.tobytes()


def test_sample_sharding_is_refused_in_reference_seeding(cbox):
    with _ctx(cbox, 32, 32, seed_mode=0) as ctx:
        ctx.set_first_sample(0)
        with pytest.raises(device.McptError, match="seed_mode 0"):
            ctx.set_first_sample(5)


def test_upload_with_a_cached_layout_equals_a_fresh_upload(cbox, mis):
    """SURVEY §8 f-2 (scene cache): mcpt_upload_scene_derived with the blob of mcpt_derive_scene gives the context the
    same device layout as mcpt_upload_scene - hits and rendered sums equal bit for bit - and refuses blobs that belong
    to other geometry or carry out-of-range links."""
    from tests.util import random_rays
    w, h = 96, 64
    blob = device.derive_scene(cbox.triangles, cbox.materials, cbox.nodes)
    rays = random_rays(50_000, cbox.nodes)
    kw = dict(max_depth=8, max_attempt=1 << 20, seed_mode=1, run_seed=5, samples_in_flight=4)
    with _ctx(cbox, w, h, **kw) as a:
        hits_a, closest_a = a.trace(rays)
        a.render(6)
        sums_a = a.read_sums()
    with device.Context(w, h, **kw) as b:
        b.upload_scene(cbox.triangles, cbox.materials, cbox.nodes, derived=blob)
        b.set_camera(cbox.camera)
        hits_b, closest_b = b.trace(rays)
        b.render(6)
        assert hit_mismatches(hits_a, hits_b) == 0 and np.array_equal(closest_a, closest_b)
        assert b.read_sums().tobytes() == sums_a.tobytes()
        # other geometry (a node box moved by one ulp): the content hash no longer matches
        moved = cbox.nodes.copy()
        moved["bbmax"][5][0] = np.nextafter(moved["bbmax"][5][0], np.float32(np.inf))
        with pytest.raises(device.McptError, match="stale"):
            b.upload_scene(cbox.triangles, cbox.materials, moved, derived=blob)
        with pytest.raises(device.McptError, match="another scene"):
            b.upload_scene(mis.triangles, mis.materials, mis.nodes, derived=blob)
        # a link of the fast tree bent out of range (header | DLeaf[n] | DNode[2n-1] | DNode[n-1]: left is word 12 of a node)
        n = len(cbox.triangles)
        bent = bytearray(blob)
        off = 80 + 96 * n + 64 * (2 * n - 1) + 64 * 7 + 48
        bent[off:off + 4] = np.int32(n + 12345).tobytes()
        with pytest.raises(device.McptError, match="out of range|twice|reach"):
            b.upload_scene(cbox.triangles, cbox.materials, cbox.nodes, derived=bytes(bent))
        with pytest.raises(device.McptError, match="too small|version"):
            b.upload_scene(cbox.triangles, cbox.materials, cbox.nodes, derived=b"garbage" * 20)
        # the refused uploads left the context as it was
        b.reset_accumulation()
        b.render(6)
        assert b.read_sums().tobytes() == sums_a.tobytes()


def test_sort_sizes_estimated_from_an_earlier_camera_do_not_lose_rays(cbox):
    """The bounce queues are sorted over an ESTIMATE of their live length taken from earlier waves.  After a camera that
    sees nothing (every path dies at once) the estimates are far too small for the real camera: items beyond the sorted
    prefix must then walk in queue order - same sums as a context that never saw the first camera."""
    w = h = 512
    kw = dict(max_depth=8, max_attempt=1 << 20, seed_mode=1, run_seed=9, samples_in_flight=2)
    away = cbox.camera.copy()
    away["direction"][0][:3] = -away["direction"][0][:3]          # looking away from the box
    with _ctx(cbox, w, h, **kw) as a:
        a.set_camera(away)
        a.render(6)                                               # three waves: their (empty) live counts are polled
        assert a.read_sums()[6].sum() == 0
        a.set_camera(cbox.camera)
        a.reset_accumulation()
        a.render(4)
        got = a.read_sums()
        rays = a.stats()["rays_traced"]
    with _ctx(cbox, w, h, **kw) as b:
        b.render(4)
        want = b.read_sums()
    assert got.tobytes() == want.tobytes() and want[6].sum() > 0
    assert rays > 6 * w * h
