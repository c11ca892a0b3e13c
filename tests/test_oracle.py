"""CPU tests of the oracle itself: internal consistency and the committed golden vectors."""
import os

import numpy as np
import pytest

from montecarlopathtracing_b200 import types as T
from oracle import oracle_api as O

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_reduced_inverse_equals_literal_cofactor_inverse():
    rng = np.random.default_rng(0)
    n = 400_000
    m = np.zeros((n, 16), np.float32)
    m[:, [0, 1, 2, 4, 5, 6, 8, 9, 10]] = rng.normal(0, 1, (n, 9)).astype(np.float32) * \
        np.float32(10.0) ** rng.integers(-3, 4, (n, 1)).astype(np.float32)
    m[rng.random(n) < 0.2, rng.integers(0, 3)] = 0.0          # zero components
    m[:, 15] = 1.0
    import ctypes
    bad = O.oracle().oracle_inverse_mismatches(m.ctypes.data_as(ctypes.c_void_p), ctypes.c_long(n))
    assert bad == 0


@pytest.mark.parametrize("name,file", [("cbox", "cbox_primary.npz"), ("mis", "mis_primary.npz")])
def test_primary_hits_match_golden(name, file, request):
    sc = request.getfixturevalue(name)
    g = np.load(os.path.join(GOLD, file))
    w, h = int(g["width"]), int(g["height"])
    rays = O.generate_rays(sc.camera, w, h)
    for mode in (0, 1):                                       # literal 4x4 cofactors, reduced form
        hits, closest, cnt = O.intersect(rays, sc.nodes, sc.triangles, mode=mode, count=True)
        assert np.array_equal(closest, g["closest"])
        assert np.array_equal(hits["t"].view(np.uint32), g["t_bits"])
        assert np.array_equal(hits["material_id"], g["material"])
        assert np.array_equal(hits["triangle_id"], g["raw_id"])
        assert np.array_equal(hits["normal"].view(np.uint32), g["normal_bits"])
        assert np.array_equal(hits["point"].view(np.uint32), g["point_bits"])
        assert (int(cnt[0]), int(cnt[1])) == (int(g["nodes_fetched"]), int(g["tris_fetched"]))


def test_render_matches_golden(cbox):
    g = np.load(os.path.join(GOLD, "cbox_render.npz"))
    for mode in (0, 1):
        r = O.render(cbox.camera, 24, 24, cbox.nodes, cbox.triangles, cbox.materials, 8, 5, 6, seed_mode=mode, run_seed=11)
        assert np.array_equal(r["hist"].view(np.uint32), g["hist_bits_%d" % mode])
        assert np.array_equal(r["counts"], g["counts_%d" % mode])
        assert r["rays_traced"] == int(g["rays_%d" % mode])


def test_lamp_wins_coplanar_ties(cbox):
    """with the luminaire emitted first and the canonical LBVH the lamp is visible."""
    rays = O.generate_rays(cbox.camera, 128, 128)
    hits, closest = O.intersect(rays, cbox.nodes, cbox.triangles)
    light = np.flatnonzero(cbox.materials["type"] == T.LIGHT)
    assert np.isin(hits["material_id"][closest >= 0], light).sum() > 20


def test_stage_loop_equals_fused_render(cbox):
    """The oracle's per-kernel functions chained like OpenCL::update() == its pixel-outer render()."""
    w = h = 20
    n, depth, cap, frames = w * h, 6, 3, 4
    seeds = O.init_seeds(n, 5)
    hist = np.zeros((n, 4), np.float32)
    counts = np.zeros(n, np.int32)
    for _ in range(frames):
        rays = O.generate_rays(cbox.camera, w, h)
        color = np.ones((n, 4), np.float32)
        hits = np.zeros(n, T.HIT)
        for _d in range(depth):
            O.intersect(rays, cbox.nodes, cbox.triangles, hits=hits)
            O.shade(cbox.materials, rays, hits, color, seeds, depth)
        O.history(color, hist, counts, cap)
    r = O.render(cbox.camera, w, h, cbox.nodes, cbox.triangles, cbox.materials, depth, cap, frames, seed_mode=0, run_seed=5)
    assert np.array_equal(r["counts"], counts) and r["hist"].tobytes() == hist.tobytes()
    assert counts.max() <= cap


def test_history_skips_zero_samples_and_caps():
    new = np.array([[0, 0, 0, 0], [2, 4, 6, 0], [1, 1, 1, 0]], np.float32)
    hist = np.array([[5, 5, 5, 0], [0, 0, 0, 0], [3, 3, 3, 0]], np.float32)
    cnt = np.array([1, 0, 2], np.int32)
    O.history(new, hist, cnt, 2)
    assert cnt.tolist() == [1, 1, 2]                       # zero sample skipped; third pixel already at the cap
    assert new.tolist() == [[5, 5, 5, 0], [2, 4, 6, 0], [3, 3, 3, 0]]
    assert hist[1].tolist() == [2, 4, 6, 0]


def test_epo_restatement_close_to_the_reference_cpu_epo():
    """kernels/EPO.cl restated in the oracle vs BVH::TEST::EPO, the reference's CPU version compiled into oracle/_ref
    (bvhtest.cpp:120-281).  The CPU version adds every contribution into one float and walks breadth-first, the
    kernel sums per triangle depth-first: same quantity, equal to about 1e-3."""
    from oracle import oracle_api as O
    from montecarlopathtracing_b200 import scenes
    if O.ref() is None or not hasattr(O.ref(), "ref_epo"):
        pytest.skip("oracle/_ref not built")
    sc = scenes.load("veach_mis")
    epo, area = O.epo_per_leaf(sc.nodes, sc.triangles)
    assert (area > 0).all() and (epo >= 0).all()
    assert O.epo_total(epo, area) == pytest.approx(O.ref_epo(sc.nodes, sc.triangles), rel=2e-3)
