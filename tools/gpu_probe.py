"""Quick device-side timing of the traversal kernels (development aid; bench.py is the contract)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from montecarlopathtracing_b200 import device, scenes, types as T
from tests.util import random_rays


def time_trace(ctx, d_rays, n, d_hits, d_closest, trav, iters=5):
    stream = torch.cuda.ExternalStream(ctx.stream)
    ctx.trace_device(d_rays.data_ptr(), n, d_hits.data_ptr(), d_closest.data_ptr(), trav)
    ctx.synchronize()
    best = 1e9
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        ctx.trace_device(d_rays.data_ptr(), n, d_hits.data_ptr(), d_closest.data_ptr(), trav)
        e1.record(stream)
        e1.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def main():
    names = sys.argv[1:] or ["cbox", "veach_mis"]
    for name in names:
        sc = scenes.load(name)
        W = H = 2048
        n = W * H
        with device.Context(W, H, 8, 16) as ctx:
            ctx.upload_scene(sc.triangles, sc.materials, sc.nodes)
            ctx.set_camera(sc.camera)
            ctx.generate_rays()
            prim = ctx.read_rays()
            rnd = random_rays(n, sc.nodes)
            d_hits = torch.empty(n * 48, dtype=torch.uint8, device="cuda")
            d_closest = torch.empty(n, dtype=torch.int32, device="cuda")
            for label, rays in (("coherent", prim), ("random", rnd)):
                d_rays = torch.from_numpy(rays.view(np.uint8)).cuda()
                res = {}
                for trav, tn in ((0, "reference"), (3, "wide8"), (1, "fast")):
                    ms = time_trace(ctx, d_rays, n, d_hits, d_closest, trav)
                    st = ctx.trace_count(d_rays.data_ptr(), n, trav)
                    res[tn] = d_closest.cpu().numpy().copy() if False else None
                    ctx.trace_device(d_rays.data_ptr(), n, d_hits.data_ptr(), d_closest.data_ptr(), trav)
                    ctx.synchronize()
                    res[tn] = (d_hits.cpu().numpy().copy(), d_closest.cpu().numpy().copy())
                    print("%s %s %s: %.3f ms  %.1f Mrays/s  nodes/ray %.2f leaves/ray %.2f tris/ray %.2f slow %d replayed %d" % (
                        name, label, tn, ms, n / ms / 1e3, st["nodes_fetched"] / n, st["leaves_fetched"] / n, st["tris_fetched"] / n,
                        st["slow_rays"], st["replayed_rays"]), flush=True)
                h0 = res["reference"][0].view(T.HIT); h1 = res["fast"][0].view(T.HIT)
                bad = int((res["reference"][1] != res["fast"][1]).sum())
                badt = int((h0["t"].view(np.uint32) != h1["t"].view(np.uint32)).sum())
                print("   fast vs reference: %d id mismatches, %d t mismatches of %d" % (bad, badt, n), flush=True)
        # render timing
        for (w, h, spp, sif, mode) in ((1024, 1024, 16, 4, 1), (1024, 1024, 8, 0, 0)):
            with device.Context(w, h, 8, 1 << 20, seed_mode=mode, samples_in_flight=sif) as ctx:
                ctx.upload_scene(sc.triangles, sc.materials, sc.nodes)
                ctx.set_camera(sc.camera)
                ctx.render(2)
                ctx.reset_stats()
                t = time.time()
                ctx.render(spp)
                dt = time.time() - t
                st = ctx.stats()
                print("%s render %dx%d %d spp mode %d: %.1f ms  %.1f Mrays/s %.1f Mpaths/s (%.2f rays/path)" % (
                    name, w, h, spp, mode, dt * 1e3, st["rays_traced"] / dt / 1e6, st["paths"] / dt / 1e6,
                    st["rays_traced"] / st["paths"]), flush=True)


if __name__ == "__main__":
    main()
