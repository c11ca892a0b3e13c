#!/usr/bin/env python3
"""Recover Wavefront OBJ geometry from the Maya-2019 binary scenes the reference ships.

The reference repository ignores ``*.obj`` (its .gitignore) so the Cornell box and Veach-MIS
geometry only survive inside ``Scene/cbox/cbox.mb`` and ``Scene/veach_mis/mis.mb``.  This tool
walks the IFF container (64-bit chunk variant), pulls vertex/edge/face tables out of every
``DMSH`` group's ``MESH`` chunk, reads the material -> shape connections from the ``CWFL``
chunks and writes an OBJ whose ``usemtl`` names are exactly the names in the scene's ``.mtl``.

Format notes:
  group  = tag 'FOR8' + 4 B + u64 size (BE) + 4-char form type, children follow, no padding
  leaf   = tag(4) + 4 B + u64 size + payload padded to 8 B
  MESH   = "o\\0", flag byte, u32 nFloats, f32 xyz..., u32 nEdgeInts, (v0|hard, v1)...,
           u32 nFaceInts, edge refs (id & 0x0FFFFFFF; bit31 = reversed; bit30 = last of face)

Mesh emission order: Maya lists nodes alphabetically; ``--first`` moves the named meshes to the
front.  For cbox the luminaire must precede the ceiling: the lamp quad is exactly coplanar with
two ceiling faces and the reference's closest-hit rule is "first visited wins ties"
(objdef.h:213), so with the stable Morton sort the lower triangle index wins.

Usage: mb2obj.py in.mb out.obj --mtllib cbox.mtl [--first cbox_luminaire_mesh]
"""
import argparse
import os
import re
import struct
import sys

GROUP_TAGS = (b"FOR8", b"LIS8", b"CAT8", b"PRO8")


def iter_chunks(d, off, end):
    while off < end:
        tag = d[off:off + 4]
        size = struct.unpack(">Q", d[off + 8:off + 16])[0]
        if tag in GROUP_TAGS:
            yield tag, d[off + 16:off + 20], off + 20, off + 16 + size
            off += 16 + size
        else:
            yield tag, None, off + 16, off + 16 + size
            off += 16 + ((size + 7) & ~7)


def cstr(d, off):
    e = d.index(b"\0", off)
    return d[off:e].decode("latin1"), e + 1


def parse_mesh(payload):
    assert payload[:2] == b"o\0", payload[:8]
    p = 3  # "o\0" + flag byte
    nf = struct.unpack(">I", payload[p:p + 4])[0]; p += 4
    verts = struct.unpack(">%df" % nf, payload[p:p + 4 * nf]); p += 4 * nf
    ne = struct.unpack(">I", payload[p:p + 4])[0]; p += 4
    edges = struct.unpack(">%dI" % ne, payload[p:p + 4 * ne]); p += 4 * ne
    nfi = struct.unpack(">I", payload[p:p + 4])[0]; p += 4
    frefs = struct.unpack(">%dI" % nfi, payload[p:p + 4 * nfi]); p += 4 * nfi
    ev = [(edges[2 * i] & 0x7FFFFFFF, edges[2 * i + 1] & 0x7FFFFFFF) for i in range(ne // 2)]
    faces, cur = [], []
    for r in frefs:
        eid = r & 0x0FFFFFFF
        a, b = ev[eid]
        if r & 0x80000000:
            a, b = b, a
        cur.append((a, b))
        if r & 0x40000000:
            # consistency: each edge's end is the next edge's start
            for i in range(len(cur)):
                assert cur[i][1] == cur[(i + 1) % len(cur)][0], "face edge loop does not chain"
            faces.append([e[0] for e in cur])
            cur = []
    assert not cur
    pts = [verts[3 * i:3 * i + 3] for i in range(nf // 3)]
    return pts, faces


def recover(path):
    d = open(path, "rb").read()
    assert d[:4] == b"FOR8" and d[16:20] == b"Maya"
    meshes = []      # (shape_name, transform_name, pts, faces)
    binds = {}       # shape_name -> material name
    for tag, form, s, e in iter_chunks(d, 20, len(d)):
        if form == b"DMSH":
            shape = parent = None
            geo = None
            for t, f, s2, e2 in iter_chunks(d, s, e):
                if t == b"CREA":
                    off = s2 + 1
                    shape, off = cstr(d, off)
                    parent, off = cstr(d, off)
                elif t == b"MESH" and d[s2:s2 + 2] == b"o\0":
                    geo = parse_mesh(d[s2:e2])
            if geo is not None:
                meshes.append((shape, parent, geo[0], geo[1]))
    # material -> shape bindings: "<mat>.mwc\0<shape>.iog.og[0].gco" inside CWFL chunks (the enclosing
    # group nesting varies between files, so scan the byte stream for the pattern instead).
    for m in re.finditer(rb"\0([^\0]+)\.mwc\0([^\0]+?)\.iog[^\0]*\0", d):
        binds[m.group(2).decode("latin1")] = m.group(1).decode("latin1")
    return meshes, binds


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("mb")
    ap.add_argument("obj")
    ap.add_argument("--mtllib", required=True)
    ap.add_argument("--first", action="append", default=[],
                    help="substring of a mesh name to emit first (repeatable)")
    a = ap.parse_args()
    meshes, binds = recover(a.mb)
    # usemtl must name a material of the .mtl exactly: the Maya file says "mis:plate_3" where the
    # course's mis.mtl says "mi:plate_3", so names are matched on the part after the namespace colon.
    mtl_names = []
    mtl_path = os.path.join(os.path.dirname(os.path.abspath(a.obj)), a.mtllib)
    if os.path.exists(mtl_path):
        mtl_names = [l.split(None, 1)[1].strip() for l in open(mtl_path) if l.startswith("newmtl")]
    def resolve(name):
        if name is None or not mtl_names or name in mtl_names:
            return name
        tail = name.split(":")[-1]
        cands = [m for m in mtl_names if m.split(":")[-1] == tail]
        if not cands:
            print("warning: material %s not in %s" % (name, a.mtllib), file=sys.stderr)
            return name
        return cands[0]
    def rank(m):
        for i, s in enumerate(a.first):
            if s in m[0] or s in (m[1] or ""):
                return i
        return len(a.first)
    order = sorted(range(len(meshes)), key=lambda i: (rank(meshes[i]), i))
    nv = nt = 0
    with open(a.obj, "w") as f:
        f.write("# recovered from %s by tools/mb2obj.py\n" % a.mb.split("/")[-1])
        f.write("mtllib %s\n" % a.mtllib)
        base = 1
        for i in order:
            shape, parent, pts, faces = meshes[i]
            key = shape.split("|")[-1]
            mat = binds.get(key) or binds.get(shape)
            if mat is None:
                cands = [v for k, v in binds.items() if k.split("|")[-1] == key]
                mat = cands[0] if cands else None
            mat = resolve(mat)
            if mat is None:
                print("warning: no material bound to", shape, file=sys.stderr)
            f.write("g %s\n" % key)
            if mat is not None:
                f.write("usemtl %s\n" % mat)
            for p in pts:
                f.write("v %.9g %.9g %.9g\n" % p)
            for fc in faces:
                f.write("f " + " ".join(str(base + v) for v in fc) + "\n")
                nt += len(fc) - 2
            base += len(pts)
            nv += len(pts)
    print("%s: %d meshes, %d verts, %d triangles after triangulation" % (a.obj, len(meshes), nv, nt))


if __name__ == "__main__":
    main()
