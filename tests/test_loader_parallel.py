"""Parallel OBJ/MTL loader: mcpt_host_load_object_mt must return the serial loader's bytes, which
tests/test_host.py pins to the reference's own loadObject (tinyobj, oracle/_ref), for any thread count."""
import os

import numpy as np
import pytest

from montecarlopathtracing_b200 import host

ASSETS = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "assets")


def _same(a, b):
    return (a.triangles.tobytes() == b.triangles.tobytes() and a.materials.tobytes() == b.materials.tobytes()
            and a.material_ids.tobytes() == b.material_ids.tobytes())


@pytest.mark.parametrize("name,obj", [("cbox", "cbox.obj"), ("veach_mis", "mis.obj"), ("diningroom", "diningroom.obj")])
def test_shipped_scenes(name, obj):
    d = os.path.join(ASSETS, name) + "/"
    if not os.path.exists(d + obj):
        pytest.skip("asset not generated")
    ref = host.load_object(d, obj)
    for th in (0, 2, 3, 7, 16):
        assert _same(ref, host.load_object(d, obj, threads=th)), th


def _fuzz_obj(rng, nfaces, eol):
    """Polygon soup with everything the line grammar allows: v / f (3..7 corners, absolute, negative and v/vt/vn
    indices), usemtl switches incl. unknown names, several mtllib lines, comments, blank lines, trailing blanks."""
    lines = ["# fuzz", "mtllib missing.mtl a.mtl b.mtl", ""]
    nv = 0
    mats = ["red", "glass", "lamp", "nosuch", "steel"]
    for f in range(nfaces):
        k = int(rng.integers(3, 8))
        c = rng.uniform(-50, 50, 3)
        ang = np.sort(rng.uniform(0, 2 * np.pi, k))
        rad = rng.uniform(0.2, 3.0, k)                      # star-shaped: concave corners exercise the ear clipping
        u, w = rng.normal(size=3), rng.normal(size=3)
        for a, r in zip(ang, rad):
            p = c + r * (np.cos(a) * u + np.sin(a) * w)
            lines.append("v %.6f %.6f %.6f%s" % (p[0], p[1], p[2], "  \t" if rng.random() < 0.1 else ""))
        nv += k
        if rng.random() < 0.15:
            lines.append("usemtl " + mats[int(rng.integers(len(mats)))])
        if rng.random() < 0.02:
            lines.append("mtllib b.mtl")
        if rng.random() < 0.05:
            lines.append("#" + "x" * int(rng.integers(0, 40)))
        style = rng.random()
        if style < 0.4:
            idx = ["%d" % (nv - k + 1 + j) for j in range(k)]
        elif style < 0.8:
            idx = ["%d" % (-(k - j)) for j in range(k)]
        else:
            idx = ["%d/%d/%d" % (nv - k + 1 + j, j + 1, 1) for j in range(k)]
        lines.append("f " + " ".join(idx))
        if rng.random() < 0.03 and nv > 20:                 # a face re-using old vertices
            lines.append("f %d %d %d %d" % tuple(int(x) for x in rng.integers(1, nv + 1, 4)))
    return eol.join(lines) + (eol if rng.random() < 0.5 else "")


MTL_A = """newmtl red
Kd 0.7 0.1 0.1
newmtl glass
Ni 1.5
Kd 1 1 1
newmtl lamp
Ka 10 9 8
"""
MTL_B = """newmtl steel
Ns 80
Ks 0.8 0.8 0.8
Kd 0.1 0.1 0.1
newmtl red
Kd 0 1 0
"""


@pytest.mark.parametrize("eol", ["\n", "\r\n", "\r"])
@pytest.mark.parametrize("nfaces", [1, 40, 6000])
def test_fuzzed_files(tmp_path, eol, nfaces):
    rng = np.random.default_rng(nfaces * 7 + len(eol))
    (tmp_path / "a.mtl").write_text(MTL_A)
    (tmp_path / "b.mtl").write_text(MTL_B)
    (tmp_path / "s.obj").write_bytes(_fuzz_obj(rng, nfaces, eol).encode())
    d = str(tmp_path) + "/"
    ref = host.load_object(d, "s.obj")
    assert len(ref.triangles) >= nfaces
    if nfaces > 1:
        assert len(set(ref.material_ids.tolist())) > 1
    for th in (0, 2, 5, 13):
        assert _same(ref, host.load_object(d, "s.obj", threads=th)), th


def test_forward_reference_fails_like_the_serial_loader(tmp_path):
    (tmp_path / "s.obj").write_text("v 0 0 0\nv 1 0 0\nf 1 2 3\nv 0 1 0\n" + "v 1 1 1\n" * 30000 + "f 1 2 3\n")
    d = str(tmp_path) + "/"
    with pytest.raises(RuntimeError, match="out of range"):
        host.load_object(d, "s.obj")
    with pytest.raises(RuntimeError, match="out of range"):
        host.load_object(d, "s.obj", threads=4)


def test_missing_file():
    with pytest.raises(RuntimeError):
        host.load_object("/nonexistent/", "x.obj", threads=4)
