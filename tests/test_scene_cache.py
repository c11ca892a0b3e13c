"""the binary scene cache.  The reference re-parses the OBJ and rebuilds the BVH on every start
(thirdpartywrapper.cpp:25-99, scenebuild.cpp:50-101); the cache must hand back exactly the bytes a fresh load produces,
notice every change of the inputs, and never accept a damaged file."""
import os
import shutil

import numpy as np
import pytest

from montecarlopathtracing_b200 import device, host, scenes


def _copy_scene(name, tmp_path):
    d, obj, _, _ = scenes.SCENES[name]
    dst = tmp_path / d
    shutil.copytree(os.path.join(scenes.ASSETS, d), dst)
    for f in os.listdir(dst):
        if f.endswith(".mcptcache"):
            os.remove(dst / f)
    return str(dst) + os.sep, obj


@pytest.mark.parametrize("name,bvhtype", [("cbox", "hlbvh"), ("veach_mis", "treelet")])
def test_cache_hit_returns_the_bytes_of_a_fresh_load(name, bvhtype, tmp_path):
    directory, obj = _copy_scene(name, tmp_path)
    fresh = host.load_object(directory, obj)
    nodes = host.build_bvh(fresh.triangles, bvhtype)
    derived = device.derive_scene(fresh.triangles, fresh.materials, nodes)
    key = host.scene_key(directory, obj, bvhtype)
    path = host.cache_path(directory, obj, bvhtype)
    assert path == directory + obj + "." + bvhtype + ".mcptcache"
    assert host.cache_load(path, key) is None and "no cache file" in host.cache_miss_reason()
    host.cache_store(path, key, fresh, nodes, derived)
    got, got_nodes, got_derived = host.cache_load(path, key)
    assert got.triangles.tobytes() == fresh.triangles.tobytes()
    assert got.materials.tobytes() == fresh.materials.tobytes()
    assert got.material_ids.tobytes() == fresh.material_ids.tobytes()
    assert got_nodes.tobytes() == nodes.tobytes()
    assert got_derived == derived
    # ... and of a fresh parallel load + a fresh derivation (the layout builder is deterministic across thread counts)
    again = host.load_object(directory, obj, threads=0)
    assert again.triangles.tobytes() == got.triangles.tobytes()
    assert device.derive_scene(again.triangles, again.materials, host.build_bvh(again.triangles, bvhtype)) == derived
    assert not [f for f in os.listdir(directory) if ".tmp." in f]          # written under a temporary name, renamed


def test_key_follows_the_content_of_obj_mtl_and_bvhtype(tmp_path):
    directory, obj = _copy_scene("cbox", tmp_path)
    k0 = host.scene_key(directory, obj, "hlbvh")
    assert k0 == host.scene_key(directory, obj, "hlbvh")
    assert k0 != host.scene_key(directory, obj, "treelet")
    assert k0 != host.scene_key(directory, obj, "hlbvh", with_normals=False)
    mtl = [f for f in os.listdir(directory) if f.endswith(".mtl")]
    assert mtl
    with open(directory + mtl[0], "a") as f:
        f.write("\n# touched\n")
    k1 = host.scene_key(directory, obj, "hlbvh")
    assert k1 != k0                                       # a material edit invalidates the cache
    with open(directory + obj, "a") as f:
        f.write("\nv 0 0 0\n")
    assert host.scene_key(directory, obj, "hlbvh") not in (k0, k1)
    with pytest.raises(RuntimeError):
        host.scene_key(directory, "missing.obj", "hlbvh")


def test_stale_truncated_and_corrupt_files_are_misses(tmp_path):
    directory, obj = _copy_scene("veach_mis", tmp_path)
    sc = host.load_object(directory, obj)
    nodes = host.build_bvh(sc.triangles)
    key = host.scene_key(directory, obj)
    path = host.cache_path(directory, obj)
    host.cache_store(path, key, sc, nodes, b"\x01\x02\x03")
    assert host.cache_load(path, key) is not None
    assert host.cache_load(path, (key[0] ^ 1, key[1])) is None and "stale" in host.cache_miss_reason()
    blob = open(path, "rb").read()
    with open(path, "wb") as f:
        f.write(blob[:-1])
    assert host.cache_load(path, key) is None and "size" in host.cache_miss_reason()
    flipped = bytearray(blob)
    flipped[len(blob) // 2] ^= 0x40                       # one bit inside the triangle / node payload
    with open(path, "wb") as f:
        f.write(bytes(flipped))
    assert host.cache_load(path, key) is None and "checksum" in host.cache_miss_reason()
    with open(path, "wb") as f:
        f.write(b"not a cache file at all")
    assert host.cache_load(path, key) is None
    with open(path, "wb") as f:
        f.write(blob)
    assert host.cache_load(path, key) is not None


def test_cache_dir_override(tmp_path, monkeypatch):
    monkeypatch.setenv("MCPT_CACHE_DIR", str(tmp_path))
    assert host.cache_path("/scenes/cbox/", "cbox.obj", "hlbvh") == str(tmp_path) + "/cbox.obj.hlbvh.mcptcache"


def test_derived_blob_is_checked_before_use(cbox):
    """mcpt_derive_scene needs no GPU; the blob names the geometry it came from."""
    blob = device.derive_scene(cbox.triangles, cbox.materials, cbox.nodes)
    n = len(cbox.triangles)
    assert len(blob) == 80 + 96 * n + 64 * (2 * n - 1) + 64 * (n - 1)
    assert blob[:8] == b"MCPTDRV1"
    with pytest.raises(device.McptError):                  # malformed input is refused like mcpt_upload_scene refuses it
        bad = cbox.nodes.copy()
        bad["left"][0] = len(bad) + 5
        device.derive_scene(cbox.triangles, cbox.materials, bad)
