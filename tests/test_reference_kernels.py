"""Pins the oracle to the reference's OWN kernel sources.  oracle/_ref/libmcpt_refcl.so is kernels/*.cl and the
device half of objdef.h compiled as C++ by oracle/build_ref.sh (syntactic adaptations listed there, OpenCL built-ins
from oracle/clc_shim.hpp); every buffer the kernels define must equal the restatement's bit for bit.
Not compared: Ray.ratio - shade.cl copies an uninitialised `Ray newRay` into rays[id] and never sets it."""
import numpy as np
import pytest

from montecarlopathtracing_b200 import scenes, types as T
from oracle import oracle_api as O
from tests.util import adversarial_rays, random_rays

pytestmark = pytest.mark.skipif(O.refcl() is None, reason="oracle/_ref/libmcpt_refcl.so not built (needs /root/reference once)")

RAY_FIELDS = ("origin", "state", "direction", "pixel")
CASES = [("cbox", (96, 96), 8), ("veach_mis", (128, 72), 16), ("diningroom", (96, 54), 16)]


def _rays_equal(a, b):
    return all(a[f].tobytes() == b[f].tobytes() for f in RAY_FIELDS)


@pytest.mark.parametrize("name,res,depth", CASES)
def test_generate_rays(name, res, depth):
    sc = scenes.load(name)
    a, b = O.generate_rays(sc.camera, *res), O.refcl_generate_rays(sc.camera, *res)
    assert a.tobytes() == b.tobytes()
    cam = sc.camera.copy()
    cam["camera_type"] = 1                                  # orthographic branch, rayGenerator.cl:23-28
    assert O.generate_rays(cam, 40, 30).tobytes() == O.refcl_generate_rays(cam, 40, 30).tobytes()


@pytest.mark.parametrize("name,res,depth", CASES)
def test_frame_loop_every_buffer(name, res, depth):
    """OpenCL::update (OpenCLApp.cpp:57-82) for three frames: rays, hits, colour, seeds after every bounce, history and
    counts after every frame, the display raster at the end."""
    sc = scenes.load(name)
    w, h = res
    n = w * h
    seeds_a = O.init_seeds(n, 11)
    seeds_b = seeds_a.copy()
    hist_a, hist_b = np.zeros((n, 4), np.float32), np.zeros((n, 4), np.float32)
    cnt_a, cnt_b = np.zeros(n, np.int32), np.zeros(n, np.int32)
    kinds = set()
    for frame in range(3):
        ra, rb = O.generate_rays(sc.camera, w, h), O.refcl_generate_rays(sc.camera, w, h)
        ca, cb = np.ones((n, 4), np.float32), np.ones((n, 4), np.float32)
        ha, hb = np.zeros(n, T.HIT), np.zeros(n, T.HIT)
        for d in range(depth):
            live = (ra["state"].view(np.uint32) & 0xFF000000) == 0
            O.intersect(ra, sc.nodes, sc.triangles, hits=ha)
            O.refcl_intersect(rb, sc.nodes, sc.triangles, hits=hb)
            assert ha[live].tobytes() == hb[live].tobytes(), (frame, d)        # incl. the raw triangleID field
            hit = live & (ha["t"] < np.float32(3e38))
            kinds.update(np.unique(sc.materials["type"][ha["material_id"][hit]]).tolist())
            O.shade(sc.materials, ra, ha, ca, seeds_a, depth)
            O.refcl_shade(sc.materials, rb, hb, cb, seeds_b, depth)
            assert _rays_equal(ra, rb), (frame, d)
            assert ca.tobytes() == cb.tobytes() and seeds_a.tobytes() == seeds_b.tobytes(), (frame, d)
        O.history(ca, hist_a, cnt_a, 2)                      # cap of 2: the third frame exercises `>= MAX_ATTEMPT`
        O.refcl_history(cb, hist_b, cnt_b, w, h, 2)
        assert ca.tobytes() == cb.tobytes() and hist_a.tobytes() == hist_b.tobytes() and np.array_equal(cnt_a, cnt_b), frame
    assert O.display(hist_a).tobytes() == O.refcl_display(hist_b, w, h).tobytes()
    assert cnt_a.sum() > 0
    if name == "cbox":
        assert kinds >= {1, 2, 3, 4}                         # diffuse, glossy, glass and the lamp were all shaded


@pytest.mark.parametrize("name", ["cbox", "veach_mis", "diningroom"])
def test_random_and_adversarial_rays(name):
    sc = scenes.load(name)
    for rays in (random_rays(20000, sc.nodes, seed=3), adversarial_rays(sc, 4000, seed=4)):
        ha, _ = O.intersect(rays, sc.nodes, sc.triangles)
        hb = O.refcl_intersect(rays.copy(), sc.nodes, sc.triangles)
        assert ha.tobytes() == hb.tobytes()
        assert (ha["t"] < np.float32(3e38)).any()


def test_render_entry_point_equals_the_kernel_loop():
    """oracle_render (what bench.py's CPU baseline and the GPU render tests use) against the reference kernels driven
    frame by frame with one persistent LCG stream per pixel (seed_mode 0)."""
    sc = scenes.load("cbox")
    w = h = 48
    n = w * h
    depth, attempt, frames = 8, 3, 5
    out = O.render(sc.camera, w, h, sc.nodes, sc.triangles, sc.materials, depth, attempt, frames, seed_mode=0, run_seed=21)
    seeds = O.init_seeds(n, 21)
    hist, cnt = np.zeros((n, 4), np.float32), np.zeros(n, np.int32)
    for _ in range(frames):
        rays = O.refcl_generate_rays(sc.camera, w, h)
        color = np.ones((n, 4), np.float32)
        hits = np.zeros(n, T.HIT)
        for _d in range(depth):
            O.refcl_intersect(rays, sc.nodes, sc.triangles, hits=hits)
            O.refcl_shade(sc.materials, rays, hits, color, seeds, depth)
        O.refcl_history(color, hist, cnt, w, h, attempt)
    assert out["hist"].tobytes() == hist.tobytes() and np.array_equal(out["counts"], cnt)
    # the same loop inside the harness (what bench.py times as the CPU baseline)
    seeds2 = O.init_seeds(n, 21)
    hist2, cnt2 = np.zeros((n, 4), np.float32), np.zeros(n, np.int32)
    traced = O.refcl_update(sc.camera, w, h, sc.nodes, sc.triangles, sc.materials, depth, attempt, frames, seeds2, hist2, cnt2)
    assert hist2.tobytes() == hist.tobytes() and np.array_equal(cnt2, cnt) and np.array_equal(seeds2, seeds)
    assert traced == out["rays_traced"]


@pytest.mark.parametrize("name", ["cbox", "veach_mis"])
def test_epo_kernel(name):
    """kernels/EPO.cl (BVH quality harness, SURVEY f-4): per-leaf EPO and triangle areas of the restatement equal the
    kernel's, on the LBVH and on the treelet-optimised tree; then the host sums of EPO_GPU (bvhtest.cpp:309-320)."""
    from montecarlopathtracing_b200 import host
    sc = scenes.load(name)
    for nodes in (sc.nodes, host.treelet(sc.nodes)):
        O.epo_max_polygon()
        epo, area = O.epo_per_leaf(nodes, sc.triangles)
        assert O.epo_max_polygon() <= 8            # above 8 the kernel's own point buffers overflow
        k_epo, k_area = O.refcl_epo(nodes, sc.triangles)
        assert epo.tobytes() == k_epo.tobytes() and area.tobytes() == k_area.tobytes()
        assert O.epo_total(epo, area) > 0
