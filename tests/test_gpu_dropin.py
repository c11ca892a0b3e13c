"""The drop-in claim, exercised: oracle/_ref/mcpt_dropin is integration/OpenCLApp_mcpt.cpp (the replacement bodies of
OpenCL::init / update that INTEGRATION.md shows) compiled against the reference's OWN config.cpp, thirdpartywrapper.cpp,
auxiliary.cpp, BVH/hlbvh.cpp, BVH/treeletBVH.cpp, tinyobj and stb (oracle/build_ref.sh), linked with libmcpt_cuda.so
and driven by a headless main that loops update() like openglapp.cpp:49.  Its <objname>.hdr must equal, byte for byte,
what this repository's own CLI (build/mcpt) writes and what the CPU oracle renders."""
import json
import os
import subprocess

import pytest

from montecarlopathtracing_b200 import host, scenes
from oracle import oracle_api as O

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DROPIN = os.path.join(ROOT, "oracle", "_ref", "mcpt_dropin")


def _reference_config(tmp, name, w, h, depth, attempt, bvhtype):
    """A config.json the REFERENCE's config.cpp parses (no comments; every key config.cpp:100-124 reads unconditionally)."""
    _, obj, _, _ = scenes.SCENES[name]
    cam = dict(host.CAMERAS[name])
    cam["resolution"] = [w, h]
    entry = {"bvhtype": bvhtype, "width": w, "height": h, "platform": "nvidia", "directory": scenes.scene_dir(name),
             "objname": obj, "maxdepth": depth, "attempt": attempt, "raygenerator": "./kernels/rayGenerator.cl",
             "intersect": "./kernels/intersect.cl", "shade": "./kernels/shade.cl", "camera": cam, "opencl": True}
    with open(os.path.join(tmp, "config.json"), "w") as f:
        json.dump({"config": [entry], "configid": 0}, f)
    return obj


@pytest.mark.skipif(not os.path.exists(DROPIN), reason="oracle/_ref/mcpt_dropin not built (oracle/build_ref.sh needs /root/reference)")
@pytest.mark.parametrize("bvhtype,batch", [("hlbvh", 1), ("hlbvh", 8), ("treelet", 3)])
def test_dropin_binary_writes_the_reference_output_file(cbox, tmp_path, bvhtype, batch):
    w = h = 192
    depth, attempt, seed = 8, 5, 21
    tmp = str(tmp_path)
    obj = _reference_config(tmp, "cbox", w, h, depth, attempt, bvhtype)
    env = dict(os.environ, MCPT_SEED=str(seed), MCPT_DROPIN_BATCH=str(batch))
    r = subprocess.run([DROPIN], cwd=tmp, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "Finished Attempting" in r.stdout                      # colorout.cpp:65
    info = json.loads(r.stdout.strip().splitlines()[-1])
    assert info["update_calls"] == attempt + 2                    # attempt+1 accumulating frames, then the writing call
    got = open(os.path.join(tmp, obj + ".hdr"), "rb").read()      # "<objname>.hdr" in the CWD, colorout.cpp:66
    # the oracle on the same tree type, through the same writer
    nodes = cbox.nodes if bvhtype == "hlbvh" else host.build_bvh(cbox.triangles, "treelet")
    ref = O.render(cbox.camera, w, h, nodes, cbox.triangles, cbox.materials, depth, attempt, attempt + 1, seed_mode=0, run_seed=seed)
    host.write_hdr(os.path.join(tmp, "oracle.hdr"), ref["hist"].reshape(h, w, 4))
    assert got == open(os.path.join(tmp, "oracle.hdr"), "rb").read()
    # this repository's CLI on the same config file
    exe = os.path.join(ROOT, "build", "mcpt")
    out = os.path.join(tmp, "cli")
    r2 = subprocess.run([exe, "--config", os.path.join(tmp, "config.json"), "--configid", "0", "--out", out, "--seed", str(seed)],
                        cwd=tmp, capture_output=True, text=True, timeout=300)
    assert r2.returncode == 0, r2.stderr
    assert got == open(out + ".hdr", "rb").read()
