"""Race evidence for the cross-SM protocols, product build: the same work repeated under varying occupancy must give the
same BYTES every time.  A data race in the treelet builder's arrival counters (the reference's own kernel is racy,
kernels/treeletBVH.cl:264-277), in the persistent kernels' work cursors or in the queue compaction shows up as a
run-to-run difference; none of these kernels has a legitimate source of nondeterminism in its results (queue ORDER may
differ, the sums are folded per pixel in sample order)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from montecarlopathtracing_b200 import device, host, scenes, types as T  # noqa: E402


def random_triangles(n, seed):
    rng = np.random.default_rng(seed)
    tris = np.zeros(n, T.TRIANGLE)
    c = rng.random((n, 1, 3)) * 100.0
    tris["v"][:, :, :3] = (c + rng.random((n, 3, 3)) * 2.0).astype(np.float32)
    return tris


def main():
    # treelet builder: 200 000 triangles, 20 repetitions, against the host optimiser once
    tris = random_triangles(200_000, 5)
    nodes = host.build_bvh(tris)
    want = host.treelet(nodes).tobytes()
    diff = sum(device.optimize_treelet(nodes).tobytes() != want for _ in range(20))
    print("treelet 200k triangles x 20 runs: %d differ from the host optimiser" % diff)
    assert diff == 0
    # LBVH builder
    want = nodes.tobytes()
    diff = sum(device.build_lbvh(tris).tobytes() != want for _ in range(20))
    print("lbvh 200k triangles x 20 runs: %d differ from the host build" % diff)
    assert diff == 0
    # the fast walk's tree (sah_build.cu: atomics on bins and partition cursors, order-free by construction)
    tris["material_id"] = 0
    mats = np.zeros(1, T.MATERIAL)
    mats["type"] = 1
    want = device.derive_scene(tris, mats, nodes)
    diff = sum(device.derive_scene(tris, mats, nodes, device=0) != want for _ in range(20))
    print("fast-walk tree 200k triangles x 20 runs: %d differ from the host builder" % diff)
    assert diff == 0
    sc = scenes.load("cbox")
    # wave pipeline (persistent walk + shade compaction + sort), different wave sizes = different schedules
    ref = None
    for k, inflight in enumerate([1, 2, 3, 4, 6, 8, 1, 2, 3, 4, 6, 8]):
        with device.Context(256, 256, 8, 1 << 20, seed_mode=1, run_seed=3, samples_in_flight=inflight) as c:
            c.upload_scene(sc.triangles, sc.materials, sc.nodes)
            c.set_camera(sc.camera)
            c.render(24)
            s = c.read_sums().tobytes()
        ref = ref or s
        assert s == ref, "wave render %d (samples in flight %d) differs" % (k, inflight)
    print("wave pipeline cbox 256x256 x 24 spp, 12 runs with 6 wave sizes: identical sums")
    # pixel chains (seed_mode 0) against the frame loop, 6 repetitions
    with device.Context(256, 256, 8, 1 << 20, seed_mode=0, run_seed=3) as c:
        c.upload_scene(sc.triangles, sc.materials, sc.nodes)
        c.set_camera(sc.camera)
        for _ in range(6):
            c.render(1)
        want = c.read_history()[0].tobytes()
    for k in range(6):
        with device.Context(256, 256, 8, 1 << 20, seed_mode=0, run_seed=3) as c:
            c.upload_scene(sc.triangles, sc.materials, sc.nodes)
            c.set_camera(sc.camera)
            c.render(6)
            assert c.read_history()[0].tobytes() == want, "pixel-chain run %d differs from the frame loop" % k
    print("pixel chains cbox 256x256 x 6 spp, 6 runs: identical to the frame loop")
    # merge kernel
    with device.Context(256, 256, 8, 1 << 20, seed_mode=1, run_seed=3) as a, device.Context(256, 256, 8, 1 << 20, seed_mode=1, run_seed=3) as b:
        for c in (a, b):
            c.upload_scene(sc.triangles, sc.materials, sc.nodes)
            c.set_camera(sc.camera)
        a.render(4)
        b.set_first_sample(4)
        b.render(4)
        sa, sb = a.read_sums(), b.read_sums()
        a.merge_sums([b])
        assert a.read_sums().tobytes() == (sa + sb).tobytes()
    print("merge kernel: equals the host sum")
    print("stress ok")


if __name__ == "__main__":
    main()
