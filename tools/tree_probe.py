import time, sys
sys.path.insert(0,'.')
from montecarlopathtracing_b200 import device, scenes
for name in ("cbox","diningroom"):
    sc=scenes.load(name)
    for rep in range(3):
        t=time.time(); b,ms=device.derive_scene(sc.triangles, sc.materials, sc.nodes, device=0, return_ms=True); t1=time.time()
        h=device.derive_scene(sc.triangles, sc.materials, sc.nodes); t2=time.time()
        print("%s: device-tree derive %.3f s (tree %.2f ms), host derive %.3f s, equal %s"%(name,t1-t,ms,t2-t1,b==h))
    for env in (0,1):
        import os; os.environ["MCPT_HOST_TREE"]=str(env)
        with device.Context(64,64,8,16,seed_mode=1) as c:
            c.synchronize(); t=time.time(); c.upload_scene(sc.triangles, sc.materials, sc.nodes); c.synchronize(); print("  upload_scene (MCPT_HOST_TREE=%d): %.3f s"%(env,time.time()-t))
        # second context: library's per-process cache makes it a copy only -> perturb geometry to force a derive
        sc.nodes["bbmax"][0][0]=float(sc.nodes["bbmax"][0][0])+1.0*(env+1)
