"""GPU treelet optimiser: mcpt_optimize_treelet must return the host optimiser's bytes, which the
CPU suite pins to the reference's own TreeletBVH<CPU> (tests/test_host.py, oracle/_ref)."""
import numpy as np
import pytest

from montecarlopathtracing_b200 import device, host, scenes, types as T

pytestmark = pytest.mark.gpu


def _random_tris(n, seed, spread=10.0, size=0.5):
    rng = np.random.default_rng(seed)
    tris = np.zeros(n, T.TRIANGLE)
    c = rng.uniform(-spread, spread, (n, 1, 3))
    tris["v"][:, :, :3] = (c + rng.normal(0, size, (n, 3, 3))).astype(np.float32)
    return tris


def _both(nodes):
    """(host bytes | None if the reference's SAH recursion does not terminate, device bytes | None likewise)"""
    try:
        want = host.treelet(nodes).tobytes()
    except RuntimeError as e:
        assert "does not terminate" in str(e)
        want = None
    try:
        got, ms = device.optimize_treelet(nodes, return_ms=True)
        got = got.tobytes()
    except device.McptError as e:
        assert "does not terminate" in str(e)
        got, ms = None, 0.0
    return want, got, ms


@pytest.mark.parametrize("name", ["cbox", "veach_mis", "diningroom"])
def test_scene_nodes_equal_host_optimiser(name):
    sc = scenes.load(name)
    want, got, ms = _both(sc.nodes)
    assert (want is None) == (got is None)
    assert want == got
    if want is not None:
        out = np.frombuffer(got, T.BVH_NODE)
        print("%s: %d nodes, device treelet pass %.3f ms, SAH %.3f -> %.3f" % (name, len(out), ms, host.sah(sc.nodes), host.sah(out)))


@pytest.mark.parametrize("n", [1, 2, 3, 4, 5, 6, 7, 8, 17, 100, 1000, 20_000, 200_000])
def test_random_triangles_equal_host_optimiser(n):
    compared = 0
    for seed in range(4 if n > 100 else 16):
        nodes = host.build_bvh(_random_tris(n, 1000 * n + seed))
        want, got, _ = _both(nodes)
        assert (want is None) == (got is None), (n, seed)
        assert want == got, (n, seed)
        compared += want is not None
    assert compared > 0 or n < 8      # tiny trees: the first leaf's triangle id is usually one of its ancestors


def test_first_leaf_triangle_id_decides_termination():
    """The first leaf slot takes its cost from 'node #triangleId' (treeletBVH.cpp:303): permuting the triangles so
    that this id is 0 (the root, an ancestor) must fail on both sides, any non-ancestor id must work."""
    tris = _random_tris(3000, 5)
    nodes = host.build_bvh(tris)
    first = len(tris) - 1
    t = int(nodes[first]["left"])
    anc = set()
    a = int(nodes[first]["parent"])
    while a != -1:
        anc.add(a)
        a = int(nodes[a]["parent"])
    bad = nodes.copy()
    bad[first]["left"] = bad[first]["right"] = 0
    want, got, _ = _both(bad)
    assert want is None and got is None
    ok = nodes.copy()
    free = next(i for i in range(first) if i not in anc and i != t)
    ok[first]["left"] = ok[first]["right"] = free
    want, got, _ = _both(ok)
    assert want is not None and want == got


def test_errors():
    with pytest.raises(device.McptError):
        device.optimize_treelet(np.zeros(4, T.BVH_NODE))
    nodes = host.build_bvh(_random_tris(200, 3))
    bad = nodes.copy()
    bad[5]["parent"] = 7                              # inconsistent parent link: rejected before any kernel runs
    with pytest.raises(device.McptError, match="binary tree"):
        device.optimize_treelet(bad)
    cyc = nodes.copy()
    cyc[3]["left"] = 0                                # child link back to the root
    with pytest.raises(device.McptError, match="binary tree"):
        device.optimize_treelet(cyc)
