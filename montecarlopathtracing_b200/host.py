"""ctypes binding of libmcpt_host.so — the host data-preparation layer (include/mcpt_host.h).

Mirrors the reference's L0 entry points: ThirdPartyWrapper::loadObject (thirdpartywrapper.cpp:25-99),
HLBVH<CPU> (BVH/hlbvh.cpp:92-200), Auxiliary::parseCamera (auxiliary.cpp:20-71), Config
(config.cpp:70-125) and outputPicture (thirdpartywrapper.cpp:14-23).
"""
import ctypes
import json
import os
from dataclasses import dataclass

import numpy as np

from . import types as T

_PKG_DIR = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class HostConfigStruct(ctypes.Structure):
    _fields_ = [("bvhtype", ctypes.c_char * 32), ("directory", ctypes.c_char * 512),
                ("objname", ctypes.c_char * 256), ("width", ctypes.c_int), ("height", ctypes.c_int),
                ("maxdepth", ctypes.c_int), ("attempt", ctypes.c_int), ("opencl", ctypes.c_int),
                ("testbvh", ctypes.c_int), ("testall", ctypes.c_int), ("has_camera", ctypes.c_int),
                ("res_x", ctypes.c_int), ("res_y", ctypes.c_int), ("camera", ctypes.c_uint8 * 80),
                ("nobjs", ctypes.c_int), ("objs", (ctypes.c_char * 128) * 32)]


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_PKG_DIR, "libmcpt_host.so")
        if not os.path.exists(path):
            raise RuntimeError("libmcpt_host.so is not built - run `python -c 'import __graft_entry__ as g; g.build()'`")
        L = ctypes.CDLL(path)
        L.mcpt_host_last_error.restype = ctypes.c_char_p
        L.mcpt_host_sah.restype = ctypes.c_double
        L.mcpt_host_free.argtypes = [ctypes.c_void_p]
        _LIB = L
    return _LIB


def _check(rc):
    if rc < 0:
        raise RuntimeError("mcpt_host: " + lib().mcpt_host_last_error().decode())
    return rc


@dataclass
class Scene:
    triangles: np.ndarray      # T.TRIANGLE
    materials: np.ndarray      # T.MATERIAL
    material_ids: np.ndarray   # int32 per triangle


def load_object(directory, objname, with_normals=True, threads=1):
    """threads=1: the serial loader (the pinned definition); 0 = all host threads, n = n threads: same bytes."""
    L = lib()
    tr, mt = ctypes.c_void_p(), ctypes.c_void_p()
    ids = ctypes.POINTER(ctypes.c_int)()
    nt, nm = ctypes.c_int(), ctypes.c_int()
    if threads == 1:
        _check(L.mcpt_host_load_object(directory.encode(), objname.encode(), int(with_normals),
                                       ctypes.byref(tr), ctypes.byref(nt), ctypes.byref(mt), ctypes.byref(nm),
                                       ctypes.byref(ids)))
    else:
        _check(L.mcpt_host_load_object_mt(directory.encode(), objname.encode(), int(with_normals), int(threads),
                                          ctypes.byref(tr), ctypes.byref(nt), ctypes.byref(mt), ctypes.byref(nm),
                                          ctypes.byref(ids)))
    try:
        tris = np.frombuffer(ctypes.string_at(tr, nt.value * 64), dtype=T.TRIANGLE).copy()
        mats = np.frombuffer(ctypes.string_at(mt, nm.value * 48), dtype=T.MATERIAL).copy()
        mids = np.frombuffer(ctypes.string_at(ids, nt.value * 4), dtype=np.int32).copy()
    finally:
        L.mcpt_host_free(tr); L.mcpt_host_free(mt); L.mcpt_host_free(ids)
    return Scene(tris, mats, mids)


class SceneKey(ctypes.Structure):
    _fields_ = [("h", ctypes.c_uint64 * 2)]


def scene_key(directory, objname, bvhtype="hlbvh", with_normals=True):
    """Content hash of the OBJ, of every MTL file it names, of bvhtype and of the record layouts (scene cache key)."""
    k = SceneKey()
    _check(lib().mcpt_host_scene_key(directory.encode(), objname.encode(), bvhtype.encode(), int(with_normals), ctypes.byref(k)))
    return (int(k.h[0]), int(k.h[1]))


def cache_path(directory, objname, bvhtype="hlbvh"):
    buf = ctypes.create_string_buffer(1024)
    _check(lib().mcpt_host_cache_path(directory.encode(), objname.encode(), bvhtype.encode(), buf, ctypes.c_size_t(1024)))
    return buf.value.decode()


def cache_store(path, key, scene, nodes, derived=b""):
    """scene: Scene (triangles, materials, material_ids); nodes: BVHNode[]; derived: device.derive_scene() blob."""
    tr = np.ascontiguousarray(scene.triangles, dtype=T.TRIANGLE)
    mt = np.ascontiguousarray(scene.materials, dtype=T.MATERIAL)
    ids = np.ascontiguousarray(scene.material_ids, dtype=np.int32)
    nd = np.ascontiguousarray(nodes, dtype=T.BVH_NODE)
    k = SceneKey((ctypes.c_uint64 * 2)(*key))
    blob = np.frombuffer(derived, dtype=np.uint8) if len(derived) else None
    _check(lib().mcpt_host_cache_store(path.encode(), ctypes.byref(k), tr.ctypes.data_as(ctypes.c_void_p), len(tr),
                                       mt.ctypes.data_as(ctypes.c_void_p), len(mt), ids.ctypes.data_as(ctypes.c_void_p),
                                       nd.ctypes.data_as(ctypes.c_void_p), len(nd),
                                       blob.ctypes.data_as(ctypes.c_void_p) if blob is not None else None,
                                       ctypes.c_size_t(len(derived))))


def cache_load(path, key):
    """(Scene, nodes, derived bytes) on a hit, None on a miss (absent / stale / other version / corrupt)."""
    L = lib()
    tr, mt, nd, dv = ctypes.c_void_p(), ctypes.c_void_p(), ctypes.c_void_p(), ctypes.c_void_p()
    ids = ctypes.POINTER(ctypes.c_int)()
    nt, nm, nn, nb = ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_size_t()
    k = SceneKey((ctypes.c_uint64 * 2)(*key))
    rc = _check(L.mcpt_host_cache_load(path.encode(), ctypes.byref(k), ctypes.byref(tr), ctypes.byref(nt), ctypes.byref(mt),
                                       ctypes.byref(nm), ctypes.byref(ids), ctypes.byref(nd), ctypes.byref(nn),
                                       ctypes.byref(dv), ctypes.byref(nb)))
    if rc == 1:
        return None
    try:
        def view(ptr, nbytes, dtype):       # one copy out of the malloc'ed buffer
            addr = ctypes.cast(ptr, ctypes.c_void_p).value
            return np.frombuffer((ctypes.c_char * nbytes).from_address(addr), dtype=dtype).copy() if nbytes else np.zeros(0, dtype)
        tris = view(tr, nt.value * 64, T.TRIANGLE)
        mats = view(mt, nm.value * 48, T.MATERIAL)
        mids = view(ids, nt.value * 4, np.int32)
        nodes = view(nd, nn.value * 64, T.BVH_NODE)
        derived = ctypes.string_at(dv, nb.value)
    finally:
        for p in (tr, mt, ids, nd, dv):
            L.mcpt_host_free(p)
    return Scene(tris, mats, mids), nodes, derived


def cache_miss_reason():
    return lib().mcpt_host_last_error().decode()


def build_bvh(triangles, bvhtype="hlbvh"):
    tris = np.ascontiguousarray(triangles, dtype=T.TRIANGLE)
    nodes = np.zeros(2 * len(tris) - 1, dtype=T.BVH_NODE)
    n = _check(lib().mcpt_host_build_bvh(tris.ctypes.data_as(ctypes.c_void_p), len(tris), bvhtype.encode(),
                                         nodes.ctypes.data_as(ctypes.c_void_p)))
    assert n == len(nodes)
    return nodes


def treelet(nodes):
    src = np.ascontiguousarray(nodes, dtype=T.BVH_NODE)
    out = np.zeros_like(src)
    _check(lib().mcpt_host_treelet(src.ctypes.data_as(ctypes.c_void_p), len(src), out.ctypes.data_as(ctypes.c_void_p)))
    return out


def sah(nodes):
    src = np.ascontiguousarray(nodes, dtype=T.BVH_NODE)
    return lib().mcpt_host_sah(src.ctypes.data_as(ctypes.c_void_p), len(src))


def parse_camera(cam):
    """cam: dict with position/lookat/up/fov (the config.json "camera" object) or its JSON text."""
    text = cam if isinstance(cam, str) else json.dumps(cam)
    out = np.zeros(1, dtype=T.CAMERA)
    _check(lib().mcpt_host_parse_camera(text.encode(), out.ctypes.data_as(ctypes.c_void_p)))
    return out


def load_config(path, configid=-1):
    s = HostConfigStruct()
    _check(lib().mcpt_host_load_config(path.encode(), configid, ctypes.byref(s)))
    d = {k: getattr(s, k) for k in ("width", "height", "maxdepth", "attempt", "opencl", "testbvh", "testall",
                                    "has_camera", "res_x", "res_y")}
    d.update(bvhtype=s.bvhtype.decode(), directory=s.directory.decode(), objname=s.objname.decode())
    d["camera"] = np.frombuffer(bytes(s.camera), dtype=T.CAMERA).copy()
    d["objs"] = [s.objs[i].value.decode() for i in range(s.nobjs)]
    return d


def write_hdr(path, rgba):
    img = np.ascontiguousarray(rgba, dtype=np.float32)
    h, w = img.shape[:2]
    _check(lib().mcpt_host_write_hdr(path.encode(), w, h, img.ctypes.data_as(ctypes.POINTER(ctypes.c_float))))


def write_png(path, rgba):
    img = np.ascontiguousarray(rgba, dtype=np.float32)
    h, w = img.shape[:2]
    _check(lib().mcpt_host_write_png(path.encode(), w, h, img.ctypes.data_as(ctypes.POINTER(ctypes.c_float))))


# Scene cameras of the three course scenes (Scene/README.md:20-61, config.json:20-26,47-53,75-81).
CAMERAS = {
    "cbox": {"position": [278, 273, -800], "lookat": [278, 273, -799], "up": [0, 1, 0], "fov": 39.3077},
    "veach_mis": {"position": [0, 2, 15], "lookat": [0, -2, 2.5], "up": [0, 1, 0], "fov": 28},
    "diningroom": {"position": [-0.5, 3, 5.5], "lookat": [-0.5, 2, 0], "up": [0, 1, 0], "fov": 60},
}
