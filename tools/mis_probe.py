"""NEE + MIS mode vs the reference-faithful estimator on the GPU (development aid): both converge to the same
image; prints relative error statistics and the variance ratio."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from montecarlopathtracing_b200 import device, scenes

name = sys.argv[1] if len(sys.argv) > 1 else "cbox"
W, H = (256, 256) if name == "cbox" else (320, 180)
spp_ref = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
spp_mis = int(sys.argv[3]) if len(sys.argv) > 3 else 4096
depth = 8
sc = scenes.load(name)
out = {}
for label, mode, spp in (("ref", 0, spp_ref), ("mis", 1, spp_mis)):
    with device.Context(W, H, depth, 1 << 30, seed_mode=1, run_seed=7, sampling=mode) as ctx:
        ctx.upload_scene(sc.triangles, sc.materials, sc.nodes)
        ctx.set_camera(sc.camera)
        t = time.time(); ctx.render(spp); dt = time.time() - t
        s = ctx.read_sums().reshape(7, -1)
        st = ctx.stats()
        out[label] = s
        print("%s: %d spp in %.2f s, %.1f Mrays/s, nonzero fraction %.4f" % (label, spp, dt, st["rays_traced"] / dt / 1e6, s[6].sum() / (spp * W * H)))
r, m = out["ref"], out["mis"]
img_r = r[:3] / np.maximum(r[6], 1); img_m = m[:3] / np.maximum(m[6], 1)
ok = (r[6] > 50) & (m[6] > 50) & (img_r.sum(0) > 1e-3)
rel = np.abs(img_m - img_r)[:, ok] / np.maximum(img_r[:, ok], 1e-6)
print("pixels compared %d of %d; mean rel err %.4f, median %.4f, 99%% %.4f" % (ok.sum(), W * H, rel.mean(), np.median(rel), np.quantile(rel, 0.99)))
print("mean image ref %s mis %s ratio %s" % (img_r[:, ok].mean(1), img_m[:, ok].mean(1), img_m[:, ok].mean(1) / img_r[:, ok].mean(1)))
print("count ratio mis/ref per spp: %.5f" % ((m[6].sum() / spp_mis) / (r[6].sum() / spp_ref)))
if spp_ref == spp_mis:
    print("count planes identical:", np.array_equal(r[6], m[6]), " max |diff|", np.abs(r[6] - m[6]).max(), " sum ref", r[6].sum(), "mis", m[6].sum())
    d = (m[6] - r[6]).reshape(H, W)
    print("pixels with more counts in mis: %d, fewer: %d" % ((d > 0).sum(), (d < 0).sum()))
    ys, xs = np.nonzero(d)
    print("rows histogram (8 bands):", np.histogram(ys, bins=8, range=(0, H))[0], " cols:", np.histogram(xs, bins=8, range=(0, W))[0])
    mat_ids = None
# ---- significance of the difference of the image means (over the compared pixels)
def mean_and_var(s, spp, mis):
    n = np.maximum(s[6], 1.0)
    mean = s[:3] / n
    if not mis:      # mean of the n non-zero samples: var = (E[x^2] - mean^2) / n
        var = np.maximum(s[3:6] / n - mean ** 2, 0.0) / n
    else:            # ratio of sums over N samples: delta method, denominator noise ignored (it is shared by both modes)
        N = float(spp)
        var = np.maximum(s[3:6] / N - (s[:3] / N) ** 2, 0.0) / N / (n / N) ** 2
    return mean, var
mr, vr = mean_and_var(r, spp_ref, False)
mm, vm = mean_and_var(m, spp_mis, True)
gm_r, gm_m = mr[:, ok].mean(1), mm[:, ok].mean(1)
se = np.sqrt((vr[:, ok].sum(1) + vm[:, ok].sum(1))) / ok.sum()
print("global means ref %s mis %s  diff/se %s" % (gm_r, gm_m, (gm_m - gm_r) / se))
z = (mm - mr)[:, ok] / np.sqrt(vr[:, ok] + vm[:, ok] + 1e-30)
print("per-pixel z: mean %.3f std %.3f  |z|>4: %d of %d" % (z.mean(), z.std(), (np.abs(z) > 4).sum(), z.size))
print("variance ratio ref/mis at equal spp (median over pixels): %.1f" % np.median((vr[:, ok] * spp_ref) / (vm[:, ok] * spp_mis + 1e-30)))
# ---- cleanest check: both modes are unbiased estimators of E[X] per pixel (numerator / spp); no ratio involved
def flux(s, spp):
    N = float(spp)
    mean = s[:3] / N
    var = np.maximum(s[3:6] / N - mean ** 2, 0.0) / N
    return mean, var
fr, vfr = flux(r, spp_ref); fm, vfm = flux(m, spp_mis)
good = (vfr + vfm) > 0
zz = (fm - fr)[good] / np.sqrt((vfr + vfm)[good])
print("E[X] per pixel: z mean %.4f std %.4f |z|>4: %d of %d (expected %.1f)" % (zz.mean(), zz.std(), (np.abs(zz) > 4).sum(), zz.size, zz.size * 6.3e-5))
tot_r, tot_m = fr.sum(1), fm.sum(1)
se_t = np.sqrt(vfr.sum(1) + vfm.sum(1))
print("total flux ref %s mis %s rel diff %s z %s" % (tot_r, tot_m, (tot_m - tot_r) / tot_r, (tot_m - tot_r) / se_t))
# ---- flux agreement by how often the chain reaches an emitter (deep-path pixels vs directly lit ones)
P = r[6] / spp_ref
edges = np.quantile(P[P > 0], [0, 0.25, 0.5, 0.75, 0.9, 1.0])
for a, b in zip(edges[:-1], edges[1:]):
    sel = (P >= a) & (P <= b)
    tr, tm = fr[:, sel].sum(), fm[:, sel].sum()
    se = np.sqrt(vfr[:, sel].sum() + vfm[:, sel].sum())
    print("P in [%.4f, %.4f]: %6d px  flux ref %.4f mis %.4f  ratio %.4f  z %.2f" % (a, b, sel.sum(), tr, tm, tm / max(tr, 1e-30), (tm - tr) / max(se, 1e-30)))
