"""Shared helpers for the parity tests (test infrastructure)."""
import numpy as np

from montecarlopathtracing_b200 import types as T


def random_rays(n, nodes, seed=0x5EED0001):
    """SURVEY §8d config 2 'random' set: origins uniform in the scene box, directions uniform on S^2."""
    rng = np.random.Generator(np.random.PCG64(seed))
    lo = nodes["bbmin"][0][:3].astype(np.float64)
    hi = nodes["bbmax"][0][:3].astype(np.float64)
    rays = np.zeros(n, T.RAY)
    rays["origin"] = (lo + (hi - lo) * rng.random((n, 3))).astype(np.float32)
    z = 1.0 - 2.0 * rng.random(n)
    phi = 2.0 * np.pi * rng.random(n)
    s = np.sqrt(np.maximum(0.0, 1.0 - z * z))
    d = np.stack([s * np.cos(phi), s * np.sin(phi), z], axis=1).astype(np.float32)
    rays["direction"] = d
    rays["pixel"] = np.arange(n, dtype=np.int64).astype(np.int32)
    return rays


def hits_equal(a, b, fields=("t", "material_id", "normal", "point")):
    return all(a[f].tobytes() == b[f].tobytes() for f in fields)


def hit_mismatches(a, b, fields=("t", "material_id", "normal", "point")):
    bad = np.zeros(len(a), bool)
    for f in fields:
        x = a[f].view(np.uint32).reshape(len(a), -1)
        y = b[f].view(np.uint32).reshape(len(b), -1)
        bad |= (x != y).any(axis=1)
    return int(bad.sum())


def adversarial_rays(sc, n, seed=77):
    """Origins uniform in the scene box, aimed EXACTLY (in float) at triangle vertices, edge midpoints and
    centroids, so hits land on shared edges/vertices and coplanar layers."""
    rng = np.random.Generator(np.random.PCG64(seed))
    nodes, tris = sc.nodes, sc.triangles
    lo = nodes["bbmin"][0][:3].astype(np.float64)
    hi = nodes["bbmax"][0][:3].astype(np.float64)
    t = rng.integers(0, len(tris), n)
    v = tris["v"][t][:, :, :3].astype(np.float64)          # (n, 3 verts, xyz)
    kind = rng.integers(0, 3, n)
    a = rng.integers(0, 3, n)
    idx = np.arange(n)
    target = np.where((kind == 0)[:, None], v[idx, a],
                      np.where((kind == 1)[:, None], 0.5 * (v[idx, a] + v[idx, (a + 1) % 3]), v.mean(axis=1)))
    org = (lo + (hi - lo) * rng.random((n, 3))).astype(np.float32)
    d = target.astype(np.float32) - org
    d = d / np.maximum(np.linalg.norm(d, axis=1, keepdims=True), 1e-20)
    rays = np.zeros(n, T.RAY)
    rays["origin"] = org
    rays["direction"] = d.astype(np.float32)
    rays["pixel"] = np.arange(n, dtype=np.int32)
    return rays
