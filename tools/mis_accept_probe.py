"""Image-level agreement of the NEE + MIS modes with a long reference-mode render (development aid for
tests/test_gpu_mis.py::test_image_agrees_with_a_long_reference_render): prints the statistics the test asserts and the
time-to-equal-noise ratio BASELINE.md quotes.  The MIS image is a ratio of two sums, so its variance comes from a
delete-one jackknife over independent batches (different run seeds), which sees every covariance.
usage: mis_accept_probe.py [scene] [log2 spp_ref] [log2 spp_mis_per_batch] [batches] [w h depth]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from montecarlopathtracing_b200 import device, scenes

name = sys.argv[1] if len(sys.argv) > 1 else "veach_mis"
spp_ref = 1 << (int(sys.argv[2]) if len(sys.argv) > 2 else 18)
spp_mis = 1 << (int(sys.argv[3]) if len(sys.argv) > 3 else 10)
B = int(sys.argv[4]) if len(sys.argv) > 4 else 16
W, H, depth = (int(sys.argv[5]), int(sys.argv[6]), int(sys.argv[7])) if len(sys.argv) > 7 else (160, 90, 16)
sc = scenes.load(name)


def render(mode, spp, seed):
    with device.Context(W, H, depth, 1 << 30, seed_mode=1, run_seed=seed, sampling=mode) as ctx:
        ctx.upload_scene(sc.triangles, sc.materials, sc.nodes)
        ctx.set_camera(sc.camera)
        t = time.time()
        ctx.render(spp)
        dt = time.time() - t
        return ctx.read_sums().reshape(7, -1), dt


def jackknife(mode):
    parts, secs = [], 0.0
    for k in range(B):
        s, dt = render(mode, spp_mis, 1000 + k)
        parts.append(s)
        secs += dt
    parts = np.stack(parts)
    tot = parts.sum(0)
    img = tot[:3] / np.maximum(tot[6], 1e-30)
    loo = tot[None] - parts
    img_k = loo[:, :3] / np.maximum(loo[:, 6:7], 1e-30)
    var = (B - 1) / B * ((img_k - img_k.mean(0)) ** 2).sum(0)
    return img, var, tot, secs


ref, t_ref = render(0, spp_ref, 7)
n = np.maximum(ref[6], 1.0)
img_r = ref[:3] / n
var_r = np.maximum(ref[3:6] / n - img_r ** 2, 0.0) / n
seen = ref[6] >= 100
print("%s %dx%d depth %d | reference: %d spp in %.2f s; pixels with >= 100 counted samples: %d of %d" % (name, W, H, depth, spp_ref, t_ref, seen.sum(), W * H))
rel_r = np.sqrt(var_r).mean(0) / np.maximum(img_r.mean(0), 1e-30)
for mode, label in ((0, "reference (self-check)"), (1, "NEE_MIS (raw count)"), (2, "NEE_MIS_RATIO")):
    img, var, tot, secs = jackknife(mode)
    ok = seen & (tot[6] > 0)
    z = (img - img_r)[:, ok] / np.sqrt(var_r[:, ok] + var[:, ok] + 1e-300)
    rel = np.sqrt(var).mean(0) / np.maximum(img.mean(0), 1e-30)
    print("%-24s %d x %d spp in %.2f s | z mean %+.3f std %.3f median|z| %.3f  |z|>4: %d of %d (%.2e)  max %.2f | image mean ratio %.4f | median rel sigma %.4f (reference run %.4f)" % (
        label, B, spp_mis, secs, z.mean(), z.std(), np.median(np.abs(z)), (np.abs(z) > 4).sum(), z.size, (np.abs(z) > 4).mean(), np.abs(z).max(),
        img[:, ok].mean() / img_r[:, ok].mean(), np.median(rel[ok]), np.median(rel_r[ok])))
    # seconds to 1 % relative sigma per pixel
    tt = secs * (rel[ok] / 0.01) ** 2
    print("%-24s time to 1 %% relative sigma: median pixel %.3g s, 90th percentile %.3g s" % ("", np.median(tt), np.quantile(tt, 0.9)))
    if mode == 2:
        # the denominator estimate against the raw count of the long reference run
        p_ref = ref[6] / spp_ref
        p_est = tot[6] / (B * spp_mis)
        se = np.sqrt(p_ref * (1 - p_ref) / spp_ref)
        print("%-24s P(X != 0): total estimate / reference = %.5f; per-pixel |diff| / P median %.4f" % ("", p_est.sum() / p_ref.sum(), np.median(np.abs(p_est - p_ref)[ok] / p_ref[ok])))
