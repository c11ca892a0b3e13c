"""CPU tests of the host data-preparation layer against the reference's OWN code (oracle/_ref, built
from /root/reference when present) and against the committed digests of its outputs."""
import ctypes
import hashlib
import json
import os

import numpy as np
import pytest

from montecarlopathtracing_b200 import host, scenes, types as T
from oracle import oracle_api as O

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def _ref():
    r = O.ref()
    if r is None:
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")
    return r


def test_struct_sizes_match_reference_headers():
    R = _ref()
    sizes = [T.CAMERA.itemsize, T.RAY.itemsize, T.HIT.itemsize, T.TRIANGLE.itemsize, T.MATERIAL.itemsize, T.BVH_NODE.itemsize]
    assert [R.ref_sizeof(i) for i in range(6)] == sizes == [80, 48, 48, 64, 48, 64]


@pytest.mark.parametrize("name", ["cbox", "veach_mis"])
def test_scene_and_bvh_match_reference_digests(name):
    """Our loader / normals / material classification / LBVH == the reference's (committed sha256)."""
    g = np.load(os.path.join(GOLD, "bvh_digests.npz"))
    sc = scenes.load(name)
    assert len(sc.triangles) == int(g[name + "_ntris"])
    assert sha(sc.triangles) == str(g[name + "_triangles"])
    assert sha(sc.materials) == str(g[name + "_materials"])
    assert sha(sc.nodes) == str(g[name + "_nodes"])
    assert sha(sc.camera) == str(g[name + "_camera"])


@pytest.mark.parametrize("name", ["cbox", "veach_mis"])
def test_bvh_memcmp_equal_to_reference_build(name):
    R = _ref()
    sc = scenes.load(name)
    nodes = np.zeros(len(sc.nodes), T.BVH_NODE)
    n = R.ref_hlbvh(sc.triangles.ctypes.data_as(ctypes.c_void_p), len(sc.triangles), nodes.ctypes.data_as(ctypes.c_void_p))
    assert n == len(nodes) and nodes.tobytes() == sc.nodes.tobytes()


@pytest.mark.parametrize("n", [2, 3, 17, 1000])
def test_bvh_random_triangles_equal_reference(n):
    R = _ref()
    rng = np.random.default_rng(n)
    tris = np.zeros(n, T.TRIANGLE)
    c = rng.uniform(-10, 10, (n, 1, 3))
    tris["v"][:, :, :3] = (c + rng.normal(0, 0.5, (n, 3, 3))).astype(np.float32)
    ours = host.build_bvh(tris)
    ref = np.zeros(2 * n - 1, T.BVH_NODE)
    assert R.ref_hlbvh(tris.ctypes.data_as(ctypes.c_void_p), n, ref.ctypes.data_as(ctypes.c_void_p)) == 2 * n - 1
    assert ours.tobytes() == ref.tobytes()


def _ref_treelet(R, nodes):
    out = np.zeros(len(nodes), T.BVH_NODE)
    src = np.ascontiguousarray(nodes)
    assert R.ref_treelet(src.ctypes.data_as(ctypes.c_void_p), len(src), out.ctypes.data_as(ctypes.c_void_p)) == len(src)
    return out


@pytest.mark.parametrize("name", ["cbox", "veach_mis"])
def test_treelet_memcmp_equal_to_reference(name):
    """bvhtype "treelet": our CPU optimiser == BVH::TreeletBVH<CPU> on the same LBVH, byte for byte."""
    R = _ref()
    sc = scenes.load(name)
    ours = host.treelet(sc.nodes)
    assert ours.tobytes() == _ref_treelet(R, sc.nodes).tobytes()
    assert ours.tobytes() != sc.nodes.tobytes()          # it did restructure something


@pytest.mark.parametrize("n", [2, 3, 4, 7, 8, 17, 100, 1000, 5000])
def test_treelet_random_triangles_equal_reference(n):
    R = _ref()
    rng = np.random.default_rng(100 + n)
    tris = np.zeros(n, T.TRIANGLE)
    c = rng.uniform(-10, 10, (n, 1, 3))
    tris["v"][:, :, :3] = (c + rng.normal(0, 0.5, (n, 3, 3))).astype(np.float32)
    lbvh = host.build_bvh(tris)
    # the reference's SAH pass reads node #<triangle id of the first leaf> (treeletBVH.cpp:303); it does not
    # terminate when that node is an ancestor of the first leaf, so only compare where it does
    first = int(lbvh[n - 1]["left"])
    p, anc = int(lbvh[n - 1]["parent"]), {n - 1}
    while p != -1:
        anc.add(p)
        p = int(lbvh[p]["parent"])
    if first in anc:
        with pytest.raises(RuntimeError):
            host.treelet(lbvh)
        return
    assert host.treelet(lbvh).tobytes() == _ref_treelet(R, lbvh).tobytes()
    assert host.build_bvh(tris, "treelet").tobytes() == host.treelet(lbvh).tobytes()


def test_treelet_keeps_a_valid_tree(cbox):
    nd = host.treelet(cbox.nodes)
    n = len(cbox.triangles)
    seen = np.zeros(len(nd), bool)
    stack = [0]
    while stack:
        i = stack.pop()
        assert not seen[i]
        seen[i] = True
        if i < n - 1:
            for c in (int(nd[i]["left"]), int(nd[i]["right"])):
                assert int(nd[c]["parent"]) == i
                assert np.all(nd[c]["bbmin"][:3] >= nd[i]["bbmin"][:3]) and np.all(nd[c]["bbmax"][:3] <= nd[i]["bbmax"][:3])
                stack.append(c)
    assert seen.all()
    assert sorted(int(x) for x in nd[n - 1:]["left"]) == list(range(n))


def test_bvh_structure_invariants(cbox):
    nd = cbox.nodes
    n = len(cbox.triangles)
    leaf = nd["left"] == nd["right"]
    assert leaf[n - 1:].all() and not leaf[:n - 1].any()                     # internal 0..N-2, leaves N-1..2N-2
    assert sorted(nd["left"][leaf].tolist()) == list(range(n))              # every triangle exactly once
    inner = np.flatnonzero(~leaf)
    for side in ("left", "right"):
        ch = nd[side][inner]
        assert (nd["bbmin"][ch][:, :3] >= nd["bbmin"][inner][:, :3]).all()  # exact containment (monotone box test)
        assert (nd["bbmax"][ch][:, :3] <= nd["bbmax"][inner][:, :3]).all()
    assert host.sah(nd) > 0


@pytest.mark.parametrize("name", ["cbox", "veach_mis", "diningroom"])
def test_camera_matches_reference(name):
    R = _ref()
    ref = np.zeros(1, T.CAMERA)
    assert R.ref_parse_camera(json.dumps(host.CAMERAS[name]).encode(), ref.ctypes.data_as(ctypes.c_void_p)) == 0
    assert host.parse_camera(host.CAMERAS[name]).tobytes() == ref.tobytes()


def test_hdr_bytes_equal_stb(tmp_path):
    R = _ref()
    rng = np.random.default_rng(1)
    img = rng.gamma(0.5, 2.0, (37, 53, 4)).astype(np.float32)
    img[3:6, 10:20] = 0.0
    a, b = str(tmp_path / "a.hdr"), str(tmp_path / "b.hdr")
    host.write_hdr(a, img)
    assert R.ref_write_hdr(b.encode(), 53, 37, img.ctypes.data_as(ctypes.POINTER(ctypes.c_float))) == 0
    assert open(a, "rb").read() == open(b, "rb").read()


def test_png_output(tmp_path):
    from PIL import Image
    img = np.zeros((8, 16, 4), np.float32)
    img[0, :, 0] = 1.0            # row 0 = image bottom -> last PNG row
    img[7, :, 1] = 0.5
    p = str(tmp_path / "x.png")
    host.write_png(p, img)
    px = np.asarray(Image.open(p).convert("RGB"))
    assert px.shape == (8, 16, 3)
    assert (px[7, :, 0] == 255).all() and (px[0, :, 1] == round(255 * 0.5 ** (1 / 2.2))).all()


CONFIG_TEXT = """
{
  # comments start with '#', as in the reference's patched json.hpp
  "config": [
    { "bvhtype": "hlbvh", "width": 256, "height": 256, "platform": "nvidia",
      "directory": "./assets/cbox/", "objname": "cbox.obj", "maxdepth": 8, "attempt": 16,   # 16 spp
      "raygenerator": "./kernels/rayGenerator.cl", "intersect": "./kernels/intersect.cl", "shade": "./kernels/shade.cl",
      "camera": { "position": [278, 273, -800], "lookat": [278, 273, -799], "up": [0, 1, 0], "fov": 39.3077,
                  "resolution": [256, 256] },
      "opencl": true },
    { "width": 64, "height": 32, "directory": "d/", "objname": "o.obj", "maxdepth": 4, "attempt": 2, "opencl": true,
      "camera": { "position": [0, 2, 15], "lookat": [0, -2, 2.5], "up": [0, 1, 0], "fov": 28, "resolution": [64, 32] } }
  ],
  "configid": 1
}
"""


def test_config_loader(tmp_path):
    p = tmp_path / "config.json"
    p.write_text(CONFIG_TEXT)
    c = host.load_config(str(p))
    assert (c["width"], c["height"], c["maxdepth"], c["attempt"], c["bvhtype"]) == (64, 32, 4, 2, "hlbvh")
    c0 = host.load_config(str(p), 0)
    assert (c0["objname"], c0["directory"], c0["attempt"], c0["res_x"]) == ("cbox.obj", "./assets/cbox/", 16, 256)
    assert c0["camera"].tobytes() == host.parse_camera(host.CAMERAS["cbox"]).tobytes()
    ref_cfg = "/root/reference/MonteCarloPathTracing/config.json"
    if os.path.exists(ref_cfg):                      # the reference's shipped file parses too
        r = host.load_config(ref_cfg, 0)
        assert (r["width"], r["maxdepth"], r["attempt"]) == (512, 16, 2048)


def test_loader_errors():
    with pytest.raises(RuntimeError):
        host.load_object("/nonexistent/", "x.obj")
    with pytest.raises(RuntimeError):
        host.build_bvh(np.zeros(4, T.TRIANGLE), "nosuchbvh")
