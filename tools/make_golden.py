#!/usr/bin/env python3
"""Generates tests/golden/*.npz (run in the build container, where /root/reference is mounted).

What each fixture pins:
  bvh_digests.npz   sha256 of the BVHNode[] the REFERENCE's own BVH/hlbvh.cpp (compiled into
                    oracle/_ref) builds for assets/cbox and assets/veach_mis, and of the Triangle[] /
                    Material[] its loadObject produces -> pins our host layer to reference outputs.
  cbox_primary.npz  closest-hit records for the cbox camera at 48x48 (t bits, ids, materials), produced by the
                    REFERENCE'S OWN kernels (kernels/rayGenerator.cl + intersect.cl compiled as C++ into
                    oracle/_ref/libmcpt_refcl.so) and asserted equal to the oracle's; `closest` and the fetch
                    counters come from the oracle (the kernels do not report them)
  mis_primary.npz   same for veach_mis at 64x36
  cbox_render.npz   render 24x24, 6 frames, depth 8, cap 5: seed mode 0 (one LCG stream per pixel, as the reference)
                    from the reference's kernels driven like OpenCL::update and asserted equal to oracle_render;
                    seed mode 1 (order-independent samples, an extension) from the oracle
The reference ships no tests or golden vectors of its own; these are outputs of its code run here.  They travel to
the GPU box, where /root/reference does not exist.
"""
import ctypes
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracing_b200 import host, scenes, types as T   # noqa: E402
from oracle import oracle_api as O                                  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def ref_scene(name):
    R = O.ref()
    d, obj, _, _ = scenes.SCENES[name]
    tr, mt = ctypes.c_void_p(), ctypes.c_void_p()
    ids = ctypes.POINTER(ctypes.c_int)()
    nt, nm = ctypes.c_int(), ctypes.c_int()
    rc = R.ref_load_object(scenes.scene_dir(name).encode(), obj.encode(), 1, ctypes.byref(tr), ctypes.byref(nt),
                           ctypes.byref(mt), ctypes.byref(nm), ctypes.byref(ids))
    assert rc == 0
    tris = np.frombuffer(ctypes.string_at(tr, nt.value * 64), dtype=T.TRIANGLE).copy()
    mats = np.frombuffer(ctypes.string_at(mt, nm.value * 48), dtype=T.MATERIAL).copy()
    nodes = np.zeros(2 * len(tris) - 1, T.BVH_NODE)
    assert R.ref_hlbvh(tris.ctypes.data_as(ctypes.c_void_p), len(tris), nodes.ctypes.data_as(ctypes.c_void_p)) == len(nodes)
    cam = np.zeros(1, T.CAMERA)
    import json
    assert R.ref_parse_camera(json.dumps(host.CAMERAS[name]).encode(), cam.ctypes.data_as(ctypes.c_void_p)) == 0
    return tris, mats, nodes, cam


def main():
    os.makedirs(OUT, exist_ok=True)
    dig = {}
    for name in ("cbox", "veach_mis"):
        tris, mats, nodes, cam = ref_scene(name)
        dig[name + "_triangles"] = sha(tris)
        dig[name + "_materials"] = sha(mats)
        dig[name + "_nodes"] = sha(nodes)
        dig[name + "_camera"] = sha(cam)
        dig[name + "_ntris"] = len(tris)
    np.savez(os.path.join(OUT, "bvh_digests.npz"), **{k: np.array(v) for k, v in dig.items()})
    for name, (w, h) in (("cbox", (48, 48)), ("veach_mis", (64, 36))):
        sc = scenes.load(name)
        rays = O.refcl_generate_rays(sc.camera, w, h)
        assert rays.tobytes() == O.generate_rays(sc.camera, w, h).tobytes()
        hits = O.refcl_intersect(rays.copy(), sc.nodes, sc.triangles)
        o_hits, closest, cnt = O.intersect(rays, sc.nodes, sc.triangles, count=True)
        assert hits.tobytes() == o_hits.tobytes()
        np.savez_compressed(os.path.join(OUT, ("cbox" if name == "cbox" else "mis") + "_primary.npz"),
                            width=w, height=h, rays_sha=np.array(sha(rays)), closest=closest,
                            t_bits=hits["t"].view(np.uint32), material=hits["material_id"], raw_id=hits["triangle_id"],
                            normal_bits=hits["normal"].view(np.uint32), point_bits=hits["point"].view(np.uint32),
                            nodes_fetched=int(cnt[0]), tris_fetched=int(cnt[1]))
    sc = scenes.load("cbox")
    out = {}
    for mode in (0, 1):
        r = O.render(sc.camera, 24, 24, sc.nodes, sc.triangles, sc.materials, 8, 5, 6, seed_mode=mode, run_seed=11)
        if mode == 0:
            seeds = O.init_seeds(24 * 24, 11)
            hist, counts = np.zeros((24 * 24, 4), np.float32), np.zeros(24 * 24, np.int32)
            traced = O.refcl_update(sc.camera, 24, 24, sc.nodes, sc.triangles, sc.materials, 8, 5, 6, seeds, hist, counts)
            assert hist.tobytes() == r["hist"].tobytes() and np.array_equal(counts, r["counts"]) and traced == r["rays_traced"]
            r = {"hist": hist, "counts": counts, "rays_traced": traced}
        out["hist_bits_%d" % mode] = r["hist"].view(np.uint32)
        out["counts_%d" % mode] = r["counts"]
        out["rays_%d" % mode] = np.array(r["rays_traced"])
    np.savez_compressed(os.path.join(OUT, "cbox_render.npz"), **out)
    print("golden fixtures written to", OUT)


if __name__ == "__main__":
    main()
