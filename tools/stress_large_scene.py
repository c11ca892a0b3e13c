"""Ad-hoc stress run (GPU box): a multi-million-triangle random soup through the device LBVH build, the device
treelet pass, the metrics, the upload and both walks; checks device LBVH == host LBVH and FAST == reference walk."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from montecarlopathtracing_b200 import device, host, scenes, types as T
from tests.util import random_rays

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
rng = np.random.default_rng(1)
tris = np.zeros(n, T.TRIANGLE)
c = rng.uniform(-100, 100, (n, 1, 3))
tris["v"][:, :, :3] = (c + rng.normal(0, 0.3, (n, 3, 3))).astype(np.float32)
a = tris["v"][:, 1, :3] - tris["v"][:, 0, :3]
b = tris["v"][:, 2, :3] - tris["v"][:, 0, :3]
nn = np.cross(a, b)
tris["normal"] = (nn / np.linalg.norm(nn, axis=1, keepdims=True)).astype(np.float32)
tris["material_id"] = 0
t0 = time.time(); nodes, ms = device.build_lbvh(tris, return_ms=True); t1 = time.time()
print("device LBVH %d tris: %.2f ms device, %.2f s wall" % (n, ms, t1 - t0))
hn = host.build_bvh(tris); t2 = time.time()
print("host LBVH %.2f s; equal: %s" % (t2 - t1, hn.tobytes() == nodes.tobytes()))
try:
    tl, ms = device.optimize_treelet(nodes, return_ms=True)
    print("device treelet %.2f ms; SAH %.2f -> %.2f" % (ms, host.sah(nodes), host.sah(tl)))
except device.McptError as e:
    print("treelet:", e)
m = device.bvh_metrics(nodes, tris, scenes.load("cbox").camera, 256, 256)
print("metrics: sah %.3f epo %.4f lcv %.3f (ms %.2f %.2f %.2f)" % (m["sah"], m["epo"], m["lcv"], m["ms_sah"], m["ms_epo"], m["ms_lcv"]))
mats = scenes.load("cbox").materials
t3 = time.time(); blob_d, tree_ms = device.derive_scene(tris, mats, nodes, device=0, return_ms=True); t4 = time.time()
blob_h = device.derive_scene(tris, mats, nodes); t5 = time.time()
print("layout derivation: tree on the device %.2f s (tree %.1f ms), all on the host %.2f s; equal: %s" % (t4 - t3, tree_ms, t5 - t4, blob_d == blob_h))
del blob_d, blob_h
with device.Context(256, 256, 8, 16) as ctx:
    t3 = time.time(); ctx.upload_scene(tris, mats, nodes); print("upload %.2f s" % (time.time() - t3))
    rays = random_rays(1 << 20, nodes)
    h1, c1 = ctx.trace(rays, device.TRAVERSAL_FAST)
    h0, c0 = ctx.trace(rays, device.TRAVERSAL_REFERENCE)
    print("FAST vs reference walk: id mismatches %d, t mismatches %d, hit fraction %.3f" %
          (int((c1 != c0).sum()), int((h1["t"].view(np.uint32) != h0["t"].view(np.uint32)).sum()), float((c1 >= 0).mean())))
