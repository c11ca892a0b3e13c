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


# ---- the binned-SAH tree of the fast walk, built on the device (csrc/cuda/sah_build.cu) ---------------------------------
@pytest.mark.parametrize("name", ["cbox", "veach_mis", "diningroom"])
def test_device_fast_tree_equals_the_host_builder(name):
    """mcpt_derive_scene_device == mcpt_derive_scene byte for byte: what mcpt_upload_scene derives for new geometry is
    the host builder's tree (accel_build.cpp), produced by level-wise kernels and one block per small subtree."""
    from montecarlopathtracing_b200 import scenes
    sc = scenes.load(name)
    want = device.derive_scene(sc.triangles, sc.materials, sc.nodes)
    got, ms = device.derive_scene(sc.triangles, sc.materials, sc.nodes, device=0, return_ms=True)
    assert got == want
    assert 0.0 < ms < 1000.0


@pytest.mark.parametrize("n,kind", [(2, "random"), (3, "random"), (17, "random"), (1025, "random"), (5000, "flat"), (3000, "same"),
                                    (20000, "lopsided"), (200000, "random")])
def test_device_fast_tree_on_synthetic_sets(n, kind):
    """Edge cases of the builder: tiny sets, triangles in one plane, identical triangles (no axis has extent: the median
    rule, on the host for more than 1 024 of them), an exponentially spaced chain (lopsided splits down to the depth at
    which the host builder switches to medians)."""
    rng = np.random.default_rng(n)
    tris = np.zeros(n, T.TRIANGLE)
    if kind == "same":
        c = np.zeros((n, 1, 3))
    elif kind == "lopsided":
        c = np.zeros((n, 1, 3))
        c[:, 0, 0] = 2.0 ** (-(np.arange(n) % 120).astype(np.float64))
        c[:, 0, 1] = rng.random(n) * 1e-3
    else:
        c = rng.random((n, 1, 3)) * 100.0
    v = c + (rng.random((1, 3, 3)) if kind == "same" else rng.random((n, 3, 3))) * 2.0
    if kind == "flat":
        v[:, :, 1] = 5.0
    tris["v"][:, :, :3] = v.astype(np.float32)
    tris["material_id"] = 0
    mats = np.zeros(1, T.MATERIAL)
    mats["type"] = 1
    nodes = host.build_bvh(tris)
    want = device.derive_scene(tris, mats, nodes)
    got = device.derive_scene(tris, mats, nodes, device=0)
    assert got == want
