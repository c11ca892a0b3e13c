#!/usr/bin/env python3
"""Procedural STAND-IN for Scene/diningroom (the reference ships neither
diningroom.obj nor diningroom.mb, so the real geometry is unrecoverable).

Deterministic; uses the material names of the reference's Scene/diningroom/diningroom.mtl and is
framed for the reference camera (position (-0.5,3,5.5), lookat (-0.5,2,0), fov 60; config.json:75-81):
a closed room with tessellated walls, a table with legs, six chairs, finely tessellated objects on
the table (spheres, tori, a "bottle" of revolution) and two emitters under the ceiling.
About 3.7e5 triangles => ~47 MB of reference-layout BVH nodes + ~24 MB of triangles.

Usage: make_diningroom.py out.obj
"""
import sys

import numpy as np


class Mesh:
    def __init__(self):
        self.parts = []      # (material, verts (n,3), tris (m,3) 0-based local)

    def add(self, mat, v, t):
        self.parts.append((mat, np.asarray(v, np.float64), np.asarray(t, np.int64)))

    def grid(self, mat, origin, eu, ev, nu, nv):
        """Parametric rectangle origin + u*eu + v*ev, nu x nv cells, normal = eu x ev."""
        o, eu, ev = (np.asarray(a, np.float64) for a in (origin, eu, ev))
        u, v = np.meshgrid(np.linspace(0, 1, nu + 1), np.linspace(0, 1, nv + 1), indexing="ij")
        pts = o + u[..., None] * eu + v[..., None] * ev
        idx = np.arange((nu + 1) * (nv + 1)).reshape(nu + 1, nv + 1)
        a, b, c, d = idx[:-1, :-1], idx[1:, :-1], idx[1:, 1:], idx[:-1, 1:]
        tris = np.concatenate([np.stack([a, b, c], -1).reshape(-1, 3), np.stack([a, c, d], -1).reshape(-1, 3)])
        self.add(mat, pts.reshape(-1, 3), tris)

    def box(self, mat, lo, hi, n=1):
        lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
        d = hi - lo
        ex, ey, ez = np.array([d[0], 0, 0]), np.array([0, d[1], 0]), np.array([0, 0, d[2]])
        self.grid(mat, lo, ey, ex, n, n)                    # z = lo, normal -z
        self.grid(mat, lo + ez, ex, ey, n, n)               # z = hi
        self.grid(mat, lo, ex, ez, n, n)                    # y = lo
        self.grid(mat, lo + ey, ez, ex, n, n)               # y = hi
        self.grid(mat, lo, ez, ey, n, n)                    # x = lo
        self.grid(mat, lo + ex, ey, ez, n, n)               # x = hi

    def revolve(self, mat, profile, center, nseg, close=True):
        """Surface of revolution about the y axis through `center`; profile = [(radius, y), ...]."""
        prof = np.asarray(profile, np.float64)
        phi = np.linspace(0, 2 * np.pi, nseg, endpoint=False)
        r, y = prof[:, 0][:, None], prof[:, 1][:, None]
        pts = np.stack([r * np.cos(phi), np.broadcast_to(y, (len(prof), nseg)), r * np.sin(phi)], -1) + np.asarray(center)
        idx = np.arange(len(prof) * nseg).reshape(len(prof), nseg)
        a, b = idx[:-1], idx[1:]
        an, bn = np.roll(a, -1, axis=1), np.roll(b, -1, axis=1)
        tris = np.concatenate([np.stack([a, an, bn], -1).reshape(-1, 3), np.stack([a, bn, b], -1).reshape(-1, 3)])
        area_ok = np.ones(len(tris), bool)
        self.add(mat, pts.reshape(-1, 3), tris[area_ok])

    def sphere(self, mat, center, radius, nlat, nlon):
        th = np.linspace(0, np.pi, nlat + 1)[1:-1]
        prof = [(radius * np.sin(t), -radius * np.cos(t)) for t in th]
        eps = radius * 1e-3
        prof = [(eps, -radius)] + prof + [(eps, radius)]
        self.revolve(mat, prof, center, nlon)

    def torus(self, mat, center, R, r, nu, nv):
        u = np.linspace(0, 2 * np.pi, nu, endpoint=False)[:, None]
        v = np.linspace(0, 2 * np.pi, nv, endpoint=False)[None, :]
        pts = np.stack([(R + r * np.cos(v)) * np.cos(u), r * np.sin(v) + 0 * u, (R + r * np.cos(v)) * np.sin(u)], -1) + np.asarray(center)
        idx = np.arange(nu * nv).reshape(nu, nv)
        a = idx
        b = np.roll(idx, -1, axis=0)
        c = np.roll(b, -1, axis=1)
        d = np.roll(idx, -1, axis=1)
        tris = np.concatenate([np.stack([a, b, c], -1).reshape(-1, 3), np.stack([a, c, d], -1).reshape(-1, 3)])
        self.add(mat, pts.reshape(-1, 3), tris)

    def write(self, path):
        nt = 0
        with open(path, "w") as f:
            f.write("# procedural stand-in for Scene/diningroom (tools/make_diningroom.py)\nmtllib diningroom.mtl\n")
            base = 1
            for k, (mat, v, t) in enumerate(self.parts):
                f.write("g part%d\nusemtl %s\n" % (k, mat))
                np.savetxt(f, v.astype(np.float32), fmt="v %.9g %.9g %.9g")
                np.savetxt(f, t + base, fmt="f %d %d %d")
                base += len(v)
                nt += len(t)
        return nt


def build():
    m = Mesh()
    W = "scene1:white"
    # room: x [-5,4], y [0,6], z [-6,7]; inward-facing, tessellated walls
    m.grid(W, (-5, 0, -6), (0, 0, 13), (9, 0, 0), 52, 36)          # floor (normal +y)
    m.grid(W, (-5, 6, -6), (9, 0, 0), (0, 0, 13), 36, 52)          # ceiling (normal -y)
    m.grid("scene1:bottle", (-5, 0, -6), (9, 0, 0), (0, 6, 0), 36, 24)   # back wall (normal +z)
    m.grid(W, (-5, 0, 7), (0, 6, 0), (9, 0, 0), 24, 36)            # front wall, behind the camera
    m.grid(W, (-5, 0, -6), (0, 6, 0), (0, 0, 13), 24, 52)          # left wall (normal +x)
    m.grid(W, (4, 0, -6), (0, 0, 13), (0, 6, 0), 52, 24)           # right wall (normal -x)
    # table
    m.box("scene1:bottle", (-3.0, 1.45, -2.6), (2.0, 1.60, 1.4), n=16)
    for x in (-2.8, 1.65):
        for z in (-2.4, 1.05):
            m.box("scene1:silver", (x, 0.0, z), (x + 0.15, 1.45, z + 0.15), n=4)
    # six chairs: seat, back, four legs
    for (cx, cz, s) in ((-2.0, 2.0, 1), (-0.5, 2.0, 1), (1.0, 2.0, 1), (-2.0, -3.2, -1), (-0.5, -3.2, -1), (1.0, -3.2, -1)):
        m.box("scene1:lamp", (cx - 0.4, 0.85, cz - 0.4), (cx + 0.4, 0.95, cz + 0.4), n=6)
        zb = cz + 0.4 * s
        m.box("scene1:lamp", (cx - 0.4, 0.95, min(zb, zb - 0.08 * s)), (cx + 0.4, 2.0, max(zb, zb - 0.08 * s)), n=6)
        for dx in (-0.36, 0.28):
            for dz in (-0.36, 0.28):
                m.box("scene1:silver", (cx + dx, 0.0, cz + dz), (cx + dx + 0.08, 0.85, cz + dz + 0.08), n=2)
    # objects on the table (finely tessellated: these carry most of the triangles)
    m.sphere("scene1:gold", (-1.9, 2.05, -0.9), 0.45, 150, 300)
    m.sphere("scene1:silver", (0.9, 1.95, -1.3), 0.35, 130, 260)
    m.sphere("scene1:white", (-0.4, 1.85, 0.6), 0.25, 100, 200)
    m.torus("scene1:gold", (0.2, 1.70, -0.3), 0.42, 0.10, 260, 90)
    m.torus("scene1:lamp", (-2.2, 1.68, 0.5), 0.30, 0.08, 200, 70)
    bottle = [(1e-4, 1.60), (0.16, 1.60), (0.17, 1.80), (0.17, 2.20), (0.13, 2.40), (0.06, 2.55), (0.05, 2.85), (0.07, 2.88), (1e-4, 2.88)]
    prof = []
    for (r0, y0), (r1, y1) in zip(bottle[:-1], bottle[1:]):
        for k in range(24):
            prof.append((r0 + (r1 - r0) * k / 24, y0 + (y1 - y0) * k / 24))
    prof.append(bottle[-1])
    m.revolve("scene1:bottle", prof, (1.4, 0.0, 0.4), 160)
    m.revolve("scene1:silver", [(r, y) for (r, y) in prof], (-1.0, 0.0, -1.8), 160)
    # emitters under the ceiling
    m.box("scene1:light1", (-2.6, 5.55, -1.6), (-1.4, 5.65, -0.4), n=2)
    m.box("scene1:light3", (0.6, 5.55, -1.6), (1.8, 5.65, -0.4), n=2)
    m.sphere("scene1:light3", (-0.5, 4.6, -0.9), 0.18, 24, 48)
    return m


if __name__ == "__main__":
    out = sys.argv[1] if len(sys.argv) > 1 else "assets/diningroom/diningroom.obj"
    n = build().write(out)
    print("%s: %d triangles" % (out, n))
