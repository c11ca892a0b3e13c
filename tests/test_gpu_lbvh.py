"""GPU LBVH build: mcpt_build_lbvh must return the host build's bytes, which the CPU suite
pins to the reference's own HLBVH<CPU> (tests/test_host.py)."""
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


@pytest.mark.parametrize("name", ["cbox", "veach_mis", "diningroom"])
def test_scene_nodes_equal_host_build(name):
    sc = scenes.load(name)
    nodes, ms = device.build_lbvh(sc.triangles, return_ms=True)
    assert nodes.tobytes() == sc.nodes.tobytes()
    print("%s: %d triangles, device build %.3f ms" % (name, len(sc.triangles), ms))


@pytest.mark.parametrize("n", [1, 2, 3, 17, 1000, 100_000])
def test_random_triangles_equal_host_build(n):
    tris = _random_tris(n, n)
    assert device.build_lbvh(tris).tobytes() == host.build_bvh(tris).tobytes()


def test_duplicate_codes_and_flat_axis():
    """Ranges of identical Morton codes split at their midpoint (hlbvh.cpp:152-156); a flat axis gives 0/0 in
    the quantiser (pinned to cell 0); 1024 clamps to 1023."""
    tris = _random_tris(5000, 7)
    tris[1000:3000] = tris[1000]                      # 2000 identical triangles -> one long run of equal codes
    tris["v"][3000:4000, :, 2] = 1.25                 # a coplanar block
    assert device.build_lbvh(tris).tobytes() == host.build_bvh(tris).tobytes()
    flat = _random_tris(3000, 8)
    flat["v"][:, :, 1] = -3.0                          # every centroid has the same y: extent 0 on that axis
    assert device.build_lbvh(flat).tobytes() == host.build_bvh(flat).tobytes()
    same = np.repeat(_random_tris(1, 9), 257)          # all codes identical
    assert device.build_lbvh(same).tobytes() == host.build_bvh(same).tobytes()


def test_errors():
    with pytest.raises(device.McptError):
        device.build_lbvh(np.zeros(0, T.TRIANGLE))
