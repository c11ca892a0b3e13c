"""The TIMED headline workload (BASELINE configs[4]: diningroom stand-in, 1920x1080, depth 16, FAST walk — what
bench.py and the driver's BENCH/SCALE records measure) pinned to the reference's own kernels on the GPU box's host
cores (oracle/_ref/libmcpt_refcl.so = kernels/*.cl + objdef.h compiled as C++), not to another CUDA path:

  * all 2 073 600 primary hits of the product walk == kernels/intersect.cl:1-28 on the same rays, bit for bit;
  * mcpt_render in the reference's seeding (seed_mode 0) == OpenCL::update (OpenCLApp.cpp:57-82) driven over the
    reference kernels: running mean, counts and ray count, bit for bit;
  * mcpt_render with order-independent seeds (seed_mode 1, what the bench times) == the oracle's render of the same
    (pixel, sample) seeds, with several samples in flight and the wave cut in the middle.
The render checks use the 960x540 view of the same camera (the slice bench.py's cpu_baseline leg also uses) so that
the reference's kernels finish in tens of seconds; the primary-hit check is at the full 1920x1080."""
import os

import numpy as np
import pytest

from montecarlopathtracing_b200 import device, scenes
from oracle import oracle_api as O
from tests.util import hit_mismatches

pytestmark = pytest.mark.gpu

W, H, D = 1920, 1080, 16
RW, RH = 960, 540


@pytest.fixture(scope="module")
def dining():
    O.set_threads()
    if O.refcl() is not None:
        O.refcl().refcl_set_threads(os.cpu_count() or 1)
    return scenes.load("diningroom")


def _ctx(sc, w, h, **kw):
    ctx = device.Context(w, h, D, 1 << 20, **kw)
    ctx.upload_scene(sc.triangles, sc.materials, sc.nodes)
    ctx.set_camera(sc.camera)
    return ctx


def test_all_primary_hits_equal_the_reference_kernels(dining):
    assert O.refcl() is not None, "oracle/_ref/libmcpt_refcl.so missing (built by oracle/build_ref.sh, shipped prebuilt)"
    rays = O.refcl_generate_rays(dining.camera, W, H)                           # kernels/rayGenerator.cl
    ref = O.refcl_intersect(rays.copy(), dining.nodes, dining.triangles)        # kernels/intersect.cl + objdef.h:102-275
    with _ctx(dining, W, H) as ctx:
        ctx.generate_rays()
        assert ctx.read_rays().tobytes() == rays.tobytes()
        ctx.intersect(device.TRAVERSAL_FAST)
        got, closest = ctx.read_hits()
        # the persistent product kernel (what mcpt_render launches), on the same rays through the host-buffer ABI
        got_p, closest_p = ctx.trace(rays, device.TRAVERSAL_FAST)
    assert hit_mismatches(got, ref) == 0
    assert hit_mismatches(got_p, ref) == 0
    assert np.array_equal(closest, closest_p)
    hit = ref["t"] < 3e38
    assert np.array_equal(hit, closest >= 0) and hit.mean() > 0.99


def test_reference_seeded_render_equals_the_reference_frame_loop(dining):
    spp = 2
    seeds = O.init_seeds(RW * RH, 7)
    hist = np.zeros((RW * RH, 4), np.float32)
    counts = np.zeros(RW * RH, np.int32)
    traced = O.refcl_update(dining.camera, RW, RH, dining.nodes, dining.triangles, dining.materials, D, 1 << 20, spp,
                            seeds, hist, counts)
    with _ctx(dining, RW, RH, seed_mode=0, run_seed=7) as ctx:
        ctx.render(spp)                              # k_pixel_chain: the default seed_mode 0 path
        g_hist, g_counts = ctx.read_history()
        g_seeds = ctx.read_seeds()
        st = ctx.stats()
    assert np.array_equal(g_counts, counts)
    assert g_hist.view(np.uint32).tobytes() == hist.view(np.uint32).tobytes()
    assert np.array_equal(g_seeds, seeds)
    assert st["rays_traced"] == traced and st["paths"] == spp * RW * RH


def test_bench_seeded_render_equals_the_oracle(dining):
    """seed_mode 1 (the bench's mode) is this repo's extension of the reference's seeding, so its oracle is the CPU
    restatement (bit-equal to the reference kernels, tests/test_reference_kernels.py) run with the same seeds."""
    spp = 3
    ref = O.render(dining.camera, RW, RH, dining.nodes, dining.triangles, dining.materials, D, 1 << 20, spp, seed_mode=1,
                   run_seed=2024, first_sample=5)
    with _ctx(dining, RW, RH, seed_mode=1, run_seed=2024, samples_in_flight=2) as ctx:     # waves of 2 + 1 samples
        ctx.set_first_sample(5)
        ctx.render(spp)
        g_hist, g_counts = ctx.read_history()
        sums = ctx.read_sums()
        st = ctx.stats()
    assert np.array_equal(g_counts, ref["counts"])
    assert g_hist.view(np.uint32).tobytes() == ref["hist"].view(np.uint32).tobytes()
    assert st["rays_traced"] == ref["rays_traced"]
    assert np.array_equal(sums[6], ref["counts"].astype(np.float64))
    assert sums[:3].tobytes() == np.ascontiguousarray(ref["sum"].T).tobytes()
    assert sums[3:6].tobytes() == np.ascontiguousarray(ref["sumsq"].T).tobytes()


@pytest.mark.parametrize("name,w,h,depth,spp", [("cbox", 1024, 1024, 8, 2), ("mis", 1280, 720, 16, 3)])
def test_other_baseline_configs_at_full_size_equal_the_reference_frame_loop(name, w, h, depth, spp, request):
    """BASELINE configs[2] (cbox 1024x1024, depth 8) and configs[3] (veach_mis 1280x720, depth 16) at their full frame
    sizes, reference seeding: running mean, counts, seeds and ray count of mcpt_render equal OpenCL::update driven over
    the reference's own kernels, bit for bit.  veach_mis is the glossy scene (Ns up to 4000): every Phong pow of the
    render goes through cr_pow and the rescaled throughput division (mcpt_device.cuh)."""
    assert O.refcl() is not None
    O.refcl().refcl_set_threads(os.cpu_count() or 1)
    sc = request.getfixturevalue(name)
    seeds = O.init_seeds(w * h, 11)
    hist = np.zeros((w * h, 4), np.float32)
    counts = np.zeros(w * h, np.int32)
    traced = O.refcl_update(sc.camera, w, h, sc.nodes, sc.triangles, sc.materials, depth, 1 << 20, spp, seeds, hist, counts)
    with device.Context(w, h, depth, 1 << 20, seed_mode=0, run_seed=11) as ctx:
        ctx.upload_scene(sc.triangles, sc.materials, sc.nodes)
        ctx.set_camera(sc.camera)
        ctx.render(spp)
        g_hist, g_counts = ctx.read_history()
        g_seeds = ctx.read_seeds()
        st = ctx.stats()
    assert np.array_equal(g_counts, counts) and counts.sum() > 0
    assert g_hist.view(np.uint32).tobytes() == hist.view(np.uint32).tobytes()
    assert np.array_equal(g_seeds, seeds)
    assert st["rays_traced"] == traced and st["paths"] == spp * w * h
