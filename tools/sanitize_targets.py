"""Small invocations of every kernel family for the checked build (tools/checked_run.sh; also usable under compute-sanitizer, tools/sanitize.sh, where that tool is open).  Each target checks its result
against the un-instrumented expectation (host builder / a second run), so a sanitizer pass is also a correctness pass.

  treelet   treelet_build.cu: cross-SM arrival counters + __threadfence (the reference's own kernel is racy,
            kernels/treeletBVH.cl:264-277)
  lbvh      lbvh_build.cu
  fasttree  sah_build.cu (the fast walk's binned-SAH tree; level-wise kernels + one block per small subtree)
  wave      k_wave_generate / k_wave_trace_persistent / k_wave_shade / k_wave_resolve + the CUB sort of the bounce queues
  chain     k_pixel_chain (seed_mode 0) and the queue-based regeneration loop
  mis       k_wave_shade_mis / k_wave_nee_add
  merge     k_merge_sums over two contexts
  stage     the five stage kernels + mcpt_trace / mcpt_trace_noalloc
  metrics   bvh_metrics.cu (SAH / EPO / LCV)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracing_b200 import device, host, scenes, types as T  # noqa: E402


def random_triangles(n, seed=3):
    rng = np.random.default_rng(seed)
    tris = np.zeros(n, T.TRIANGLE)
    c = rng.random((n, 1, 3)) * 100.0
    tris["v"][:, :, :3] = (c + rng.random((n, 3, 3)) * 2.0).astype(np.float32)
    return tris


def ctx_for(sc, w, h, **kw):
    kw.setdefault("max_depth", 8)
    kw.setdefault("max_attempt", 1 << 20)
    c = device.Context(w, h, **kw)
    c.upload_scene(sc.triangles, sc.materials, sc.nodes)
    c.set_camera(sc.camera)
    return c


def main(target, scale):
    if target == "treelet":
        tris = random_triangles(int(20000 * scale))
        nodes = host.build_bvh(tris)
        got = device.optimize_treelet(nodes)
        assert got.tobytes() == host.treelet(nodes).tobytes()
    elif target == "fasttree":
        for n, seed in ((int(60000 * scale), 4), (900, 5), (2, 6)):
            tris = random_triangles(n, seed)
            tris["material_id"] = 0
            mats = np.zeros(1, T.MATERIAL)
            mats["type"] = 1
            nodes = host.build_bvh(tris)
            assert device.derive_scene(tris, mats, nodes, device=0) == device.derive_scene(tris, mats, nodes)
    elif target == "lbvh":
        tris = random_triangles(int(50000 * scale))
        assert device.build_lbvh(tris).tobytes() == host.build_bvh(tris).tobytes()
    elif target in ("wave", "chain", "mis", "merge", "stage"):
        sc = scenes.load("cbox")
        w = h = max(32, int(256 * min(1.0, scale)))
        if target == "wave":
            with ctx_for(sc, w, h, seed_mode=1, run_seed=3, samples_in_flight=2) as a, ctx_for(sc, w, h, seed_mode=1, run_seed=3, samples_in_flight=1) as b:
                a.render(4)
                b.render(4)
                assert a.read_sums().tobytes() == b.read_sums().tobytes()
        elif target == "chain":
            with ctx_for(sc, w, h, seed_mode=0, run_seed=3) as a, ctx_for(sc, w, h, seed_mode=0, run_seed=3) as b:
                a.render(3)                      # pixel chains
                for _ in range(3):
                    b.render(1)                  # frame loop
                assert a.read_history()[0].tobytes() == b.read_history()[0].tobytes()
            os.environ["MCPT_OPT"] = "64"        # queue-based regeneration
            with ctx_for(sc, w, h, seed_mode=0, run_seed=3) as c:
                c.render(3)
        elif target == "mis":
            with ctx_for(sc, w, h, seed_mode=1, run_seed=3, sampling=1, samples_in_flight=2) as a:
                a.render(4)
                assert np.isfinite(a.read_sums()).all()
        elif target == "merge":
            with ctx_for(sc, w, h, seed_mode=1, run_seed=3) as a, ctx_for(sc, w, h, seed_mode=1, run_seed=3) as b:
                a.render(2)
                b.set_first_sample(2)
                b.render(2)
                sa, sb = a.read_sums(), b.read_sums()
                a.merge_sums([b])
                assert a.read_sums().tobytes() == (sa + sb).tobytes()
        else:
            with ctx_for(sc, w, h, seed_mode=0, run_seed=3) as a:
                for _ in range(2):
                    a.frame()
                a.display()
                rays = np.zeros(w * h, T.RAY)
                a.generate_rays()
                rays = a.read_rays()
                h1, c1 = a.trace(rays)
                h2, c2 = a.trace_noalloc(rays[:device.TRACE_NOALLOC_MAX_RAYS])
                assert np.array_equal(c1[:len(c2)], c2)
                for trav in (device.TRAVERSAL_REFERENCE, device.TRAVERSAL_CANONICAL):
                    _, c3 = a.trace(rays, traversal=trav)
                    assert np.array_equal(c1, c3)
    elif target == "sorted":
        # coherence-sorted queues (waves of >= 2^18 paths): reference mode and NEE + MIS, including sort sizes estimated
        # under another camera (the walk's sorted-prefix guard) - each against the unsorted pipeline
        sc = scenes.load("cbox")
        w = h = 512
        kw = dict(seed_mode=1, run_seed=3, samples_in_flight=2)
        away = sc.camera.copy()
        away["direction"][0][:3] = -away["direction"][0][:3]
        for sampling in (0, 1):
            os.environ.pop("MCPT_SORT", None)
            with ctx_for(sc, w, h, sampling=sampling, **kw) as a:
                a.set_camera(away)
                a.render(6)
                a.set_camera(sc.camera)
                a.reset_accumulation()
                a.render(8)
                got = a.read_sums().tobytes()
            os.environ["MCPT_SORT"] = "0"
            with ctx_for(sc, w, h, sampling=sampling, **kw) as b:
                b.render(8)
                assert got == b.read_sums().tobytes()
        os.environ.pop("MCPT_SORT", None)
    elif target == "metrics":
        sc = scenes.load("veach_mis")
        q = device.bvh_metrics(sc.nodes, sc.triangles, sc.camera, 64, 48)
        assert q["sah"] > 0
    else:
        raise SystemExit("unknown target " + target)
    try:
        n, code, value, bound = device.check_failures()
        assert n == 0, "checked build: %d violations, first: code %d value %d bound %d" % (n, code, value, bound)
        print("target %s ok (checked build: 0 index / protocol violations)" % target)
    except device.McptError:
        print("target %s ok" % target)


if __name__ == "__main__":
    main(sys.argv[1], float(sys.argv[2]) if len(sys.argv) > 2 else 1.0)
