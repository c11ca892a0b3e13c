#!/usr/bin/env python3
"""Generates tests/golden/cbox_converged.npz: the CPU oracle's render of Scene/cbox at 20x20, depth 8, at a sample
count high enough that north_star's tolerance (mean relative error of linear radiance <= 1 %, every pixel within
4 sigma) can be TESTED rather than drowned in Monte-Carlo noise.

Why 20x20 and 2 Mi spp: 97.7 % of cbox's BSDF-sampled paths never reach the lamp and the image is the mean of the
NON-ZERO samples (history.cl:10-12), whose coefficient of variation per pixel is 1.4 (median) to 3 (90th
percentile).  A pixel therefore needs ~40 000 non-zero samples, i.e. ~2 Mi paths, before its own noise is below
1 %; at 0.3 Mpaths/s on 8 host cores that is affordable for 400 pixels (about an hour), not for 65 536.  The
reference's generator shoots ONE fixed ray through (x/W, y/H) per pixel (rayGenerator.cl:9-20, no jitter), so the
20x20 frame is exactly every 16th pixel of a 320x320 frame: a sparse, converged sample of the same image.

The oracle (oracle/mcpt_oracle.cpp) is bit-equal to the reference's own kernels compiled as C++
(tests/test_reference_kernels.py).  The run is cut into chunks with independent per-pixel LCG streams (run_seed
+ k); the sums of the chunks are one estimate.  Partial results are saved after every chunk."""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracing_b200 import scenes   # noqa: E402
from oracle import oracle_api as O               # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=20)
    ap.add_argument("--depth", type=int, default=8)
    ap.add_argument("--chunks", type=int, default=16)
    ap.add_argument("--spp-per-chunk", type=int, default=131072)
    ap.add_argument("--run-seed", type=int, default=4242)
    ap.add_argument("--threads", type=int, default=6)
    ap.add_argument("--out", default=os.path.join(ROOT, "tests", "golden", "cbox_converged.npz"))
    a = ap.parse_args()
    sc = scenes.load("cbox")
    O.set_threads(a.threads)
    n = a.size * a.size
    sums = np.zeros((7, n))
    rays = 0
    first = 0
    if os.path.exists(a.out):           # resume: chunk k always uses run_seed + k, so a continued run equals an unbroken one
        g = np.load(a.out)
        if (int(g["width"]), int(g["depth"]), int(g["run_seed"])) == (a.size, a.depth, a.run_seed) and \
                int(g["spp"]) == int(g["chunks"]) * a.spp_per_chunk:
            sums, rays, first = g["sums"].copy(), int(g["rays_traced"]), int(g["chunks"])
            print("resuming after chunk %d" % first, flush=True)
    t0 = time.time()
    for k in range(first, a.chunks):
        r = O.render(sc.camera, a.size, a.size, sc.nodes, sc.triangles, sc.materials, a.depth, 1 << 30, a.spp_per_chunk,
                     seed_mode=0, run_seed=a.run_seed + k)
        sums[:3] += r["sum"].T
        sums[3:6] += r["sumsq"].T
        sums[6] += r["counts"]
        rays += r["rays_traced"]
        np.savez_compressed(a.out, sums=sums, width=a.size, height=a.size, depth=a.depth, spp=(k + 1) * a.spp_per_chunk,
                            run_seed=a.run_seed, chunks=k + 1, rays_traced=rays)
        print("chunk %d/%d  %.0f s  %.2f Mrays/s" % (k + 1, a.chunks, time.time() - t0, rays / (time.time() - t0) / 1e6), flush=True)


if __name__ == "__main__":
    main()
