"""smoke(): one small invocation of the hot path on cuda:0, checked against the CPU oracle."""
import numpy as np


def run():
    from . import device, scenes
    from oracle import oracle_api as O      # the checker (allowed in smoke only)

    W = H = 128
    sc = scenes.load("cbox")
    with device.Context(W, H, max_depth=8, max_attempt=4, seed_mode=0, run_seed=7) as ctx:
        ctx.upload_scene(sc.triangles, sc.materials, sc.nodes)
        ctx.set_camera(sc.camera)
        ctx.generate_rays()
        rays = ctx.read_rays()
        ref_rays = O.generate_rays(sc.camera, W, H)
        assert rays.tobytes() == ref_rays.tobytes(), "primary rays differ from the oracle"
        ref_hits, ref_closest = O.intersect(ref_rays, sc.nodes, sc.triangles)
        for trav in (device.TRAVERSAL_REFERENCE, device.TRAVERSAL_FAST):
            ctx.intersect(trav)
            hits, closest = ctx.read_hits()
            assert np.array_equal(closest, ref_closest), "closest-hit ids differ (traversal %d)" % trav
            for f in ("t", "material_id", "normal", "point"):
                assert hits[f].tobytes() == ref_hits[f].tobytes(), "Hit.%s differs (traversal %d)" % (f, trav)
        ctx.reset_stats()
        ctx.render(3)
        hist, counts = ctx.read_history()
        ref = O.render(sc.camera, W, H, sc.nodes, sc.triangles, sc.materials, 8, 4, 3, seed_mode=0, run_seed=7)
        assert np.array_equal(counts, ref["counts"]), "non-zero sample counts differ"
        assert hist.tobytes() == ref["hist"].tobytes(), "accumulated image differs from the oracle"
        st = ctx.stats()
        assert st["rays_traced"] == ref["rays_traced"], (st, ref["rays_traced"])
    print("smoke ok: %d rays bit-exact, 3-spp render bit-exact, %d rays traced" % (W * H, st["rays_traced"]))
