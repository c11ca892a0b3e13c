"""north_star's image tolerance, tested as stated: "mean relative error of linear radiance at most 1 % at high spp,
every pixel within 4 sigma of its sample-variance estimate" (BASELINE config 3: Scene/cbox, depth 8, 1 B200).

What "high spp" means for this estimator: the reference's image is the mean of the NON-ZERO samples
(history.cl:10-12), 97.7 % of cbox's BSDF-sampled paths are zero, and the non-zero ones have a per-pixel coefficient
of variation of 1.4 (median) .. 3 (90th percentile).  At config 3's 1 024 spp a pixel holds ~23 counted samples, so
two independent renders of it differ by ~40 % — the 1 % statement is about CONVERGED pixels.  Two tests therefore:

  test_config3_*            cbox 256x256 x 1 024 spp on the GPU (order-independent seeds) against the oracle's render
                            at equal spp (per-pixel LCG streams, other seed), computed here on the host cores:
                            z-scores of every pixel/channel from sigma^2 = s_gpu^2/n_gpu + s_ref^2/n_ref.  With ~23
                            samples behind each sigma the z-scores are Student-t like, not Gaussian: P(|t_30| > 4)
                            = 4e-4, P(> 6) = 1.4e-6, so of 196 608 values ~80 may exceed 4 and none should exceed
                            7; asserted: at most 2e-3 of them beyond 4, none beyond 7 — and the unbiased global checks.
  test_converged_*          the 20x20 view (= every 16th pixel of the 320x320 frame: the generator shoots one fixed ray
                            per pixel, rayGenerator.cl:9-20) at 4 Mi + 4 Mi spp on two contexts against the committed
                            oracle render at 2 Mi spp (tests/golden/cbox_converged.npz, tools/make_golden_converged.py,
                            about an hour of host time): here sigma rests on ~10^5 samples, z is Gaussian, 1 200
                            values have an expected 0.08 beyond 4 sigma, and the assertion is north_star's own:
                            EVERY pixel within 4 sigma, mean relative error <= 1 %; for one context alone and for
                            the device-merged sums of both (the multi-GPU merge path, mcpt_merge_sums).
  test_veach_mis_*          BASELINE config 4's frame (1280x720, depth 16) at reduced spp against the oracle: block
                            means (16x16 pixels) within 4 sigma, counted-sample totals within their Poisson bound."""
import os

import numpy as np
import pytest

from montecarlopathtracing_b200 import device, sharding
from oracle import oracle_api as O

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cbox_converged.npz")


def _ctx(sc, w, h, depth, **kw):
    ctx = device.Context(w, h, depth, 1 << 30, **kw)
    ctx.upload_scene(sc.triangles, sc.materials, sc.nodes)
    ctx.set_camera(sc.camera)
    return ctx


def _oracle_sums(r):
    s = np.zeros((7, len(r["counts"])))
    s[:3], s[3:6], s[6] = r["sum"].T, r["sumsq"].T, r["counts"]
    return s


def _z(a, b, min_count):
    """|mean_a - mean_b| / sqrt(sigma_a^2 + sigma_b^2) per channel and pixel, over pixels both renders resolve."""
    both = (a[6] >= min_count) & (b[6] >= min_count)
    ma, mb = a[:3] / np.maximum(a[6], 1), b[:3] / np.maximum(b[6], 1)
    sig = np.sqrt(sharding.sigma_of_mean(a) ** 2 + sharding.sigma_of_mean(b) ** 2)
    lit = both[None, :] & (sig > 0)
    return np.abs(ma - mb)[lit] / sig[lit], ma, mb, both


def test_config3_cbox_1024spp_against_the_oracle_at_equal_spp(cbox):
    w = h = 256
    spp = 1024
    O.set_threads()
    ref = _oracle_sums(O.render(cbox.camera, w, h, cbox.nodes, cbox.triangles, cbox.materials, 8, 1 << 30, spp, seed_mode=0,
                                run_seed=101))
    with _ctx(cbox, w, h, 8, seed_mode=1, run_seed=202) as ctx:
        ctx.render(spp)
        got = ctx.read_sums()
    z, mg, mr, both = _z(got, ref, 8)                 # pixels with at least 8 counted samples in both renders (~70 %)
    assert both.mean() > 0.6 and z.size > 100_000
    assert (z > 4.0).mean() < 2e-3 and z.max() < 7.0, ((z > 4.0).sum(), z.max())
    # z^2 averages ~1.1 when the two renders estimate the same image (two oracle renders with different seeds at 96x96:
    # 1.12, 7e-4 beyond 4 sigma, max 4.75 - the small-sample sigma is itself noisy)
    assert 0.8 < np.mean(z ** 2) < 1.3
    # unbiasedness where noise averages out: frame mean of the image and the total of counted samples
    assert abs(mg[:, both].mean() / mr[:, both].mean() - 1.0) < 0.01
    assert abs(got[6].sum() - ref[6].sum()) < 4.0 * np.sqrt(got[6].sum() + ref[6].sum())


def _mre(a, b, eps):
    ma, mb = a[:3] / np.maximum(a[6], 1), b[:3] / np.maximum(b[6], 1)
    lum = mb.mean(axis=0)
    lit = lum > eps
    return float(np.mean(np.abs(ma - mb)[:, lit] / mb[:, lit])), int(lit.sum())


def test_converged_cbox_every_pixel_within_4_sigma_and_mre_below_1_percent(cbox):
    assert os.path.exists(GOLDEN), "tests/golden/cbox_converged.npz missing (tools/make_golden_converged.py)"
    g = np.load(GOLDEN)
    ref, w, h = g["sums"], int(g["width"]), int(g["height"])
    assert int(g["depth"]) == 8 and int(g["spp"]) >= (1 << 21)
    half = 1 << 22
    with _ctx(cbox, w, h, 8, seed_mode=1, run_seed=303, samples_in_flight=32768) as a, \
            _ctx(cbox, w, h, 8, seed_mode=1, run_seed=303, samples_in_flight=32768) as b:
        a.render_async(half)
        b.set_first_sample(half)
        b.render_async(half)
        a.synchronize()
        b.synchronize()
        one = a.read_sums()
        a.merge_sums([b])                           # the multi-GPU merge kernel (both contexts on this device)
        merged = a.read_sums()
    assert merged[6].sum() > 1.9 * one[6].sum()
    for label, got in (("one context, 4 Mi spp", one), ("merged, 8 Mi spp", merged)):
        z, _, _, both = _z(got, ref, 1000)
        assert both.sum() >= 0.9 * w * h, label
        assert z.max() < 4.0, (label, float(z.max()))                    # EVERY pixel and channel within 4 sigma
        assert 0.7 < np.mean(z ** 2) < 1.4, (label, float(np.mean(z ** 2)))
    # Mean relative error of linear radiance over pixels with oracle mean > eps.  39 of the 400 pixels look past the box
    # (no sample ever counted), and the dim ones are where the committed oracle render is itself unconverged: its own
    # relative sigma at 2 Mi spp averages 1.95 % over all lit pixels (17 % in the darkest), so two perfect renderers
    # would show an MRE of 1.7 % at eps = 1e-3.  eps = 0.03 (the median lit pixel is 0.045) keeps 224 pixels whose
    # expected MRE from sampling noise alone is 0.7 %: what is asserted is north_star's 1 % on those.
    mre, lit = _mre(merged, ref, 0.03)
    assert lit >= 200, lit
    assert mre <= 0.01, mre
    # and no systematic offset on ANY lit pixel class: the signed relative difference averages out (its standard error
    # over the 361 x 3 resolved values is ~0.1 %)
    ma, mb = merged[:3] / np.maximum(merged[6], 1), ref[:3] / np.maximum(ref[6], 1)
    sig = np.sqrt(sharding.sigma_of_mean(merged) ** 2 + sharding.sigma_of_mean(ref) ** 2)
    ok = ((merged[6] >= 1000) & (ref[6] >= 1000))[None, :] & (mb > 0) & (sig > 0)
    wgt = (mb[ok] / sig[ok]) ** 2                                        # inverse variance of the relative difference
    bias = float(np.sum(wgt * (ma[ok] - mb[ok]) / mb[ok]) / np.sum(wgt))
    assert abs(bias) < 4.0 / np.sqrt(np.sum(wgt)) + 1e-4, (bias, 1.0 / np.sqrt(np.sum(wgt)))


def test_veach_mis_720p_against_the_oracle(mis):
    w, h, spp, depth = 1280, 720, 32, 16      # ~1 % of veach_mis's paths reach a lamp: 8 192 paths per block give ~80 counted samples
    O.set_threads()
    ref = _oracle_sums(O.render(mis.camera, w, h, mis.nodes, mis.triangles, mis.materials, depth, 1 << 30, spp, seed_mode=0,
                                run_seed=77))
    with _ctx(mis, w, h, depth, seed_mode=1, run_seed=78) as ctx:
        ctx.render(spp)
        got = ctx.read_sums()

    def blocks(s):       # 16x16-pixel block totals of the seven planes
        return s.reshape(7, h // 16, 16, w // 16, 16).sum(axis=(2, 4)).reshape(7, -1)
    z, _, _, both = _z(blocks(got), blocks(ref), 32)
    assert both.mean() > 0.3
    assert (z > 4.0).mean() < 2e-3 and z.max() < 7.0, ((z > 4.0).sum(), z.max())
    # pooling ignores that every pixel contributes a fixed number of paths (stratification), so the pooled sigma is an
    # over-estimate wherever a block straddles an edge: only the upper bound is meaningful
    assert np.mean(z ** 2) < 1.3
    assert abs(got[6].sum() - ref[6].sum()) < 4.0 * np.sqrt(got[6].sum() + ref[6].sum())
