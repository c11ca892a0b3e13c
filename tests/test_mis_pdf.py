"""The solid-angle density of the reference's randomDirection (shade.cl:40-59) that the NEE + MIS mode uses
(random_direction_pdf in csrc/cuda/mcpt_device.cuh), checked against draws of the reference's own discrete
process: u = r2 / 32768, direction ~ (sqrt(u) cos phi, sqrt(u) sin phi, 1 - u) normalised."""
import numpy as np


def pdf(c):
    c = np.minimum(c, 1.0)
    v = 2.0 * c / (c + np.sqrt(4.0 - 3.0 * c * c))
    u = 1.0 - v
    D = 1.0 - u + u * u
    return np.where(c > 0, D * np.sqrt(D) / (np.pi * (1.0 + u)), 0.0)


def test_density_matches_the_discrete_process():
    rng = np.random.default_rng(5)
    u = rng.integers(0, 32768, 4_000_000) / 32768.0          # float(random()) / 32768.0f
    cos_theta = (1.0 - u) / np.sqrt(1.0 - u + u * u)
    edges = np.linspace(0.0, 1.0, 41)
    hist, _ = np.histogram(cos_theta, bins=edges)
    # P(cos in [a, b]) = 2 pi * integral of pdf(c) dc (phi is uniform)
    c = np.linspace(0.0, 1.0, 400_001)
    cdf = np.concatenate([[0.0], np.cumsum(0.5 * (pdf(c[1:]) + pdf(c[:-1])) * np.diff(c))]) * 2.0 * np.pi
    expect = np.interp(edges[1:], c, cdf) - np.interp(edges[:-1], c, cdf)
    assert abs(cdf[-1] - 1.0) < 1e-5                          # it is a density on the hemisphere
    got = hist / hist.sum()
    assert np.max(np.abs(got - expect) / expect) < 0.02
    assert abs(pdf(np.array([1.0]))[0] - 1.0 / np.pi) < 1e-12 and abs(pdf(np.array([1e-9]))[0] - 0.5 / np.pi) < 1e-6


def test_mean_cosine():
    """SURVEY §7.4 quotes E[cos theta] = 0.5494 for the process."""
    c = np.linspace(0.0, 1.0, 400_001)
    f = pdf(c) * c * 2.0 * np.pi
    assert abs(np.sum(0.5 * (f[1:] + f[:-1]) * np.diff(c)) - 0.5494) < 5e-4


def test_glossy_lobe_acceptance_probability_matches_the_rejection_loop():
    """shade.cl:130-132 redraws randomDirection(reflect) until the direction is above the surface; the NEE + MIS mode
    divides the lobe's density by the loop's acceptance probability, tabulated by the library against
    theta = angle(reflect, normal).  Checked against draws of the discrete process itself."""
    from montecarlopathtracing_b200 import device
    tab = device.phong_accept_table()
    assert tab.shape == (1025,) and tab[0] == 1.0 and abs(tab[-1] - 0.5) < 1e-4
    assert np.all(np.diff(tab) <= 1e-7)                       # monotone: the more the lobe tilts, the more is cut off
    rng = np.random.default_rng(9)
    n = 2_000_000
    for theta_deg in (5.0, 30.0, 60.0, 85.0):
        theta = np.radians(theta_deg)
        u = rng.integers(0, 32768, n) / 32768.0
        phi = 2.0 * np.pi * rng.integers(0, 32768, n) / 32768.0
        s = np.sqrt(u)
        # lobe frame: axis r = z; the surface normal tilted by theta in the x-z plane
        d = np.stack([s * np.cos(phi), s * np.sin(phi), 1.0 - u], axis=1)
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        normal = np.array([np.sin(theta), 0.0, np.cos(theta)])
        accepted = np.mean(d @ normal > 0.0)
        x = theta / (np.pi / 2) * 1024
        i = int(x)
        interp = tab[i] + (tab[i + 1] - tab[i]) * (x - i)
        assert abs(accepted - interp) < 4.0 * np.sqrt(0.25 / n) + 1e-4, (theta_deg, accepted, interp)
