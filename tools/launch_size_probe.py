"""Duration of the persistent walk kernel against the number of rays per launch (primary rays of the scene camera):
separates the kernel's fixed cost from its per-ray cost."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from montecarlopathtracing_b200 import device, scenes

name = sys.argv[1] if len(sys.argv) > 1 else "cbox"
sc = scenes.load(name)
side = 4096
with device.Context(side, side, 1, 1) as g:
    g.upload_scene(sc.triangles, sc.materials, sc.nodes)
    g.set_camera(sc.camera)
    g.generate_rays()
    rays = torch.from_numpy(g.read_rays().view(np.uint8)).cuda()
    n_all = side * side
    perm = torch.randperm(n_all, device="cuda")
    shuffled = rays.view(n_all, 48)[perm].contiguous().view(-1)          # incoherent order, like bounce rays
    hits = torch.empty(n_all * 48, dtype=torch.uint8, device="cuda")
    closest = torch.empty(n_all, dtype=torch.int32, device="cuda")
    stream = torch.cuda.ExternalStream(g.stream)
    for label, buf in (("coherent", rays), ("shuffled", shuffled)):
        for n in (1 << 14, 1 << 16, 1 << 18, 1 << 19, 1 << 20, 1 << 21, 1 << 22, 1 << 24):
            best = 1e9
            for it in range(5):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                g.trace_device(buf.data_ptr(), n, hits.data_ptr(), closest.data_ptr(), device.TRAVERSAL_FAST)
                e1.record(stream)
                e1.synchronize()
                best = min(best, e0.elapsed_time(e1))
            print("%s %-9s n=%9d  %8.1f us  %7.1f Mrays/s" % (name, label, n, best * 1e3, n / best / 1e3))
