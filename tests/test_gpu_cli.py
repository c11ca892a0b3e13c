"""build/mcpt features either side of the hot path: the binary scene cache, checkpoint / resume of the
progressive state (colorout.cpp:23-24,49-50,63-68 keeps it on the device for the whole run) and the one-process-per-device
mode whose host-side merge must give the doubles of mcpt_merge_sums."""
import json
import os
import shutil
import subprocess

import numpy as np
import pytest

from montecarlopathtracing_b200 import device, host, scenes, sharding

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "build", "mcpt")


def _config(tmp_path, name, width, height, depth, attempt, bvhtype="hlbvh"):
    d, obj, _, _ = scenes.SCENES[name]
    scene = tmp_path / d
    if not scene.exists():
        shutil.copytree(os.path.join(scenes.ASSETS, d), scene)
        for f in os.listdir(scene):
            if f.endswith(".mcptcache"):
                os.remove(scene / f)
    cam = dict(host.CAMERAS[name], resolution=[width, height])
    cfg = {"configid": 0, "config": [{"bvhtype": bvhtype, "width": width, "height": height, "directory": str(scene) + os.sep,
                                       "objname": obj, "maxdepth": depth, "attempt": attempt, "camera": cam, "opencl": True}]}
    path = tmp_path / "config.json"
    path.write_text(json.dumps(cfg))
    return str(path), str(scene) + os.sep, obj


def _run(args, env=None):
    r = subprocess.run([EXE] + args, capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stderr + r.stdout
    assert "Finished Attempting" in r.stdout
    return json.loads(r.stdout.strip().splitlines()[-1])


def _bytes(path):
    with open(path, "rb") as f:
        return f.read()


# (treelet on veach_mis: cbox's lamp is coplanar with its ceiling and the treelet order lets the ceiling win the tie,
# SURVEY H1 - the reference-faithful image is then black whatever the materials are)
@pytest.mark.parametrize("name,bvhtype", [("cbox", "hlbvh"), ("veach_mis", "treelet")])
def test_scene_cache_miss_then_hit_same_image(tmp_path, name, bvhtype):
    cfg, directory, obj = _config(tmp_path, name, 96, 64, 8, 3, bvhtype)
    a = _run(["--config", cfg, "--out", str(tmp_path / "a"), "--seed", "9", "--no-cache"])
    assert a["scene_cache"] == "off" and not [f for f in os.listdir(directory) if f.endswith(".mcptcache")]
    b = _run(["--config", cfg, "--out", str(tmp_path / "b"), "--seed", "9"])
    assert b["scene_cache"] == "miss" and os.path.exists(host.cache_path(directory, obj, bvhtype))
    c = _run(["--config", cfg, "--out", str(tmp_path / "c"), "--seed", "9"])
    assert c["scene_cache"] == "hit"
    assert _bytes(str(tmp_path / "a.hdr")) == _bytes(str(tmp_path / "b.hdr")) == _bytes(str(tmp_path / "c.hdr"))
    # the file holds what a fresh load + build + derivation produce
    sc = host.load_object(directory, obj)
    nodes = host.build_bvh(sc.triangles, bvhtype)
    got, got_nodes, derived = host.cache_load(host.cache_path(directory, obj, bvhtype), host.scene_key(directory, obj, bvhtype))
    assert got.triangles.tobytes() == sc.triangles.tobytes() and got_nodes.tobytes() == nodes.tobytes()
    assert derived == device.derive_scene(sc.triangles, sc.materials, nodes)
    # an edited material file is noticed: miss again, and the image changes
    mtl = [f for f in os.listdir(directory) if f.endswith(".mtl")][0]
    text = open(directory + mtl).read().replace("Kd 0.", "Kd 0.3")      # every material's red reflectance
    open(directory + mtl, "w").write(text)
    d = _run(["--config", cfg, "--out", str(tmp_path / "d"), "--seed", "9"])
    assert d["scene_cache"] == "miss"
    assert _bytes(str(tmp_path / "d.hdr")) != _bytes(str(tmp_path / "a.hdr"))
    assert len(set(_bytes(str(tmp_path / "a.hdr"))[80:])) > 8          # not a black frame


@pytest.mark.parametrize("seeding", [[], ["--order-independent-seeds"]])
def test_checkpointed_and_resumed_runs_write_the_same_file(tmp_path, seeding):
    cfg, _, _ = _config(tmp_path, "cbox", 80, 60, 8, 10)           # attempt 10 -> 11 frames
    base = ["--config", cfg, "--seed", "4"] + seeding
    _run(base + ["--out", str(tmp_path / "full")])
    ck = str(tmp_path / "state.ckp")
    _run(base + ["--out", str(tmp_path / "cut"), "--checkpoint", ck, "--checkpoint-every", "4"])     # 4 + 4 + 3
    assert _bytes(str(tmp_path / "cut.hdr")) == _bytes(str(tmp_path / "full.hdr"))
    # an interrupted run (8 of the 11 frames), then resume to 11 from the file
    _run(base + ["--out", str(tmp_path / "part"), "--stop-after", "8", "--checkpoint", ck, "--checkpoint-every", "8"])
    assert _bytes(str(tmp_path / "part.hdr")) != _bytes(str(tmp_path / "full.hdr"))
    _run(base + ["--out", str(tmp_path / "resumed"), "--resume", ck])
    assert _bytes(str(tmp_path / "resumed.hdr")) == _bytes(str(tmp_path / "full.hdr"))
    r = subprocess.run([EXE] + base + ["--out", str(tmp_path / "x"), "--seed", "5", "--resume", ck], capture_output=True, text=True)
    assert r.returncode != 0 and "another size, depth, seed" in r.stderr


def test_one_process_per_device_equals_the_in_process_merge(tmp_path, cbox):
    """--procs 2 (here both workers on the one visible device): each renders its block of sample indices and the parent
    adds the sum planes in rank order - the same image as two contexts merged by mcpt_merge_sums."""
    w, h, attempt = 96, 64, 8                                    # 9 frames: 5 + 4
    cfg, _, _ = _config(tmp_path, "cbox", w, h, 8, attempt)
    env = dict(os.environ, CUDA_VISIBLE_DEVICES=os.environ.get("CUDA_VISIBLE_DEVICES", "0").split(",")[0] * 1)
    env["CUDA_VISIBLE_DEVICES"] = env["CUDA_VISIBLE_DEVICES"] + "," + env["CUDA_VISIBLE_DEVICES"]
    for extra, state in ((["--no-cache"], "off"), ([], "miss"), ([], "hit")):
        out = str(tmp_path / ("p_" + state))
        line = _run(["--config", cfg, "--out", out, "--seed", "12", "--procs", "2"] + extra, env=env)
        assert line["procs"] == 2 and line["scene_cache"] == state
    kw = dict(max_depth=8, max_attempt=attempt, seed_mode=1, run_seed=12)
    with device.Context(w, h, **kw) as a, device.Context(w, h, **kw) as b:
        for c in (a, b):
            c.upload_scene(cbox.triangles, cbox.materials, cbox.nodes)
            c.set_camera(cbox.camera)
        a.render(5)
        b.set_first_sample(5)
        b.render(4)
        a.merge_sums([b])
        img = device.image_from_sums(a.read_sums(), w, h)
        host.write_hdr(str(tmp_path / "ref.hdr"), img)
        # --overlap-init: the workers initialise CUDA one after the other and pull blocks of 32 frames from a shared
        # counter; these 9 frames are one block, so one worker renders them all and the other contributes zeros
        a.reset_accumulation()
        a.render(9)
        host.write_hdr(str(tmp_path / "ref_one.hdr"), device.image_from_sums(a.read_sums(), w, h))
    ref = _bytes(str(tmp_path / "ref.hdr"))
    for state in ("off", "miss", "hit"):
        assert _bytes(str(tmp_path / ("p_" + state + ".hdr"))) == ref, state
    line = _run(["--config", cfg, "--out", str(tmp_path / "stag"), "--seed", "12", "--procs", "2", "--overlap-init"], env=env)
    assert line["procs"] == 2 and line["first_gpu_ready_s"] <= line["all_gpus_ready_s"]
    assert _bytes(str(tmp_path / "stag.hdr")) == _bytes(str(tmp_path / "ref_one.hdr"))


def test_a_refused_cached_layout_falls_back_to_a_fresh_derivation(tmp_path):
    """A cache file whose arrays are current but whose traversal layout was written by another builder version (here:
    a blob with a foreign header) must not stop the renderer: the layout is derived afresh and the image is the same."""
    cfg, directory, obj = _config(tmp_path, "cbox", 64, 48, 8, 2)
    _run(["--config", cfg, "--out", str(tmp_path / "a"), "--seed", "3", "--no-cache"])
    sc = host.load_object(directory, obj)
    nodes = host.build_bvh(sc.triangles)
    blob = bytearray(device.derive_scene(sc.triangles, sc.materials, nodes))
    blob[8:12] = (999).to_bytes(4, "little")                      # the blob's version field
    host.cache_store(host.cache_path(directory, obj, "hlbvh"), host.scene_key(directory, obj, "hlbvh"), sc, nodes, bytes(blob))
    r = subprocess.run([EXE, "--config", cfg, "--out", str(tmp_path / "b"), "--seed", "3"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "cached traversal layout not used" in r.stderr and '"scene_cache": "hit"' in r.stdout
    assert _bytes(str(tmp_path / "a.hdr")) == _bytes(str(tmp_path / "b.hdr"))
