#!/usr/bin/env python3
"""Upper bound of what re-ordering the bounce queue can buy the walk kernel (development probe, GPU box).

Collects REAL bounce rays of the workload (stage kernels: generate, k x (intersect, shade), several frames with
different seeds), then times the product walk (k_intersect_persistent through mcpt_trace_device) on the rays as the
pipeline would hand them over (compaction order) and after sorting them by a set of candidate keys:
  oct            direction octant (3 bits)
  mN             Morton code of the origin on a 2^N grid per axis
  mN+oct         Morton code, octant in the low bits        oct+mN: octant in the high bits
  mN+dK          Morton code + direction on a 2^K x 2^K octahedral grid
Prints Mrays/s per key and the time torch needs for the sort itself (an upper bound for a radix sort of that width)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from montecarlopathtracing_b200 import device, scenes


def spread3(v):      # 10 bits -> every third bit
    v = v & 0x3FF
    v = (v | (v << 16)) & 0x030000FF
    v = (v | (v << 8)) & 0x0300F00F
    v = (v | (v << 4)) & 0x030C30C3
    v = (v | (v << 2)) & 0x09249249
    return v


def morton(o, lo, hi, bits):
    q = ((o - lo) / (hi - lo) * (1 << bits)).clamp(0, (1 << bits) - 1).to(torch.int64)
    return spread3(q[:, 0]) | (spread3(q[:, 1]) << 1) | (spread3(q[:, 2]) << 2)


def octant(d):
    return (d[:, 0] < 0).to(torch.int64) | ((d[:, 1] < 0).to(torch.int64) << 1) | ((d[:, 2] < 0).to(torch.int64) << 2)


def octa(d, bits):
    n = d / d.abs().sum(dim=1, keepdim=True)
    u, v = n[:, 0].clone(), n[:, 1].clone()
    neg = n[:, 2] < 0
    uu = (1 - n[:, 1].abs()) * torch.sign(n[:, 0])
    vv = (1 - n[:, 0].abs()) * torch.sign(n[:, 1])
    u[neg], v[neg] = uu[neg], vv[neg]
    qu = ((u * 0.5 + 0.5) * (1 << bits)).clamp(0, (1 << bits) - 1).to(torch.int64)
    qv = ((v * 0.5 + 0.5) * (1 << bits)).clamp(0, (1 << bits) - 1).to(torch.int64)
    return (qu << bits) | qv


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "diningroom"
    bounce = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    frames = int(sys.argv[3]) if len(sys.argv) > 3 else 8
    W, H, D = 1920, 1080, 16
    sc = scenes.load(name)
    chunks = []
    with device.Context(W, H, D, 1 << 20, seed_mode=0, run_seed=1) as ctx:
        ctx.upload_scene(sc.triangles, sc.materials, sc.nodes)
        ctx.set_camera(sc.camera)
        for f in range(frames):
            ctx.reset_color()
            ctx.generate_rays()
            for _ in range(bounce):
                ctx.intersect(device.TRAVERSAL_FAST)
                ctx.shade()
            r = ctx.read_rays()
            live = (r["state"].view(np.uint32) & 0xFF000000) == 0
            chunks.append(r[live])
        rays = np.concatenate(chunks)
        n = len(rays)
        print("scene %s bounce %d: %d live rays from %d frames" % (name, bounce, n, frames))
        t = torch.from_numpy(rays.view(np.uint8).reshape(n, 48)).cuda()
        f32 = t.view(torch.float32).view(n, 12)
        o, d = f32[:, 0:3].contiguous(), f32[:, 4:7].contiguous()
        lo = torch.tensor(sc.nodes["bbmin"][0][:3].astype(np.float32), device="cuda")
        hi = torch.tensor(sc.nodes["bbmax"][0][:3].astype(np.float32), device="cuda")
        d_hits = torch.empty(n * 48, dtype=torch.uint8, device="cuda")
        d_closest = torch.empty(n, dtype=torch.int32, device="cuda")
        stream = torch.cuda.ExternalStream(ctx.stream)

        def timed(rt):
            best = 1e30
            for it in range(5):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize()
                e0.record(stream)
                ctx.trace_device(rt.data_ptr(), n, d_hits.data_ptr(), d_closest.data_ptr(), device.TRAVERSAL_FAST)
                e1.record(stream)
                e1.synchronize()
                if it:
                    best = min(best, e0.elapsed_time(e1))
            return best

        base = timed(t)
        print("%-14s %8.1f Mrays/s" % ("as queued", n / base / 1e3))
        perm = torch.randperm(n, device="cuda")
        print("%-14s %8.1f Mrays/s" % ("shuffled", n / timed(t[perm].contiguous()) / 1e3))
        keys = {"oct": octant(d)}
        if os.environ.get("PROBE_SET", "1") == "1":
            for b in (3, 4, 5, 6, 7):
                m = morton(o, lo, hi, b)
                keys["m%d" % b] = m
                keys["m%d+oct" % b] = (m << 3) | octant(d)
                keys["oct+m%d" % b] = (octant(d) << (3 * b)) | m
            for b, k in ((4, 2), (5, 2), (5, 3), (6, 3), (6, 4)):
                keys["m%d+d%d" % (b, k)] = (morton(o, lo, hi, b) << (2 * k)) | octa(d, k)
                keys["d%d+m%d" % (k, b)] = (octa(d, k) << (3 * b)) | morton(o, lo, hi, b)
        else:       # direction-major keys of 12..30 bits
            for k, b in ((3, 3), (4, 3), (4, 4), (4, 5), (5, 3), (5, 4), (5, 5), (6, 4), (6, 6), (4, 6), (5, 6), (7, 5), (8, 4), (6, 0), (8, 0)):
                keys["d%d+m%d (%d bits)" % (k, b, 2 * k + 3 * b)] = (octa(d, k) << (3 * b)) | (morton(o, lo, hi, b) if b else 0)
        for label, key in keys.items():
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            order = torch.argsort(key.to(torch.int32))
            e1.record()
            torch.cuda.synchronize()
            ms = timed(t[order].contiguous())
            print("%-14s %8.1f Mrays/s  (x%.2f; torch argsort %.2f ms, walk %.2f ms)" % (label, n / ms / 1e3, base / ms, e0.elapsed_time(e1), ms))


if __name__ == "__main__":
    main()
