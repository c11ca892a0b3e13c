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
