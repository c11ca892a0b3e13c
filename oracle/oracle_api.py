"""TEST INFRASTRUCTURE — ctypes binding of oracle/liboracle.so (CPU restatement of the reference
kernels) and oracle/_ref/libmcpt_ref.so (the reference's own host code, built by build_ref.sh).
Importable only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ORACLE = None
_REF = None

TRI = np.dtype([("v", np.float32, (3, 4)), ("normal", np.float32, (3,)), ("material_id", np.int32)])


def _vp(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(os.path.join(_HERE, "mcpt_oracle.cpp")):
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


def oracle():
    global _ORACLE
    if _ORACLE is None:
        L = ctypes.CDLL(build())
        L.oracle_sample_seed.restype = ctypes.c_uint32
        L.oracle_inverse_mismatches.restype = ctypes.c_long
        L.oracle_epo_total.restype = ctypes.c_double
        _ORACLE = L
    return _ORACLE


def ref():
    """The reference's own host code (loadObject / HLBVH / treelet / parseCamera / stb hdr)."""
    global _REF
    if _REF is None:
        so = os.path.join(_HERE, "_ref", "libmcpt_ref.so")
        if not os.path.exists(so):
            subprocess.check_call(["bash", os.path.join(_HERE, "build_ref.sh")])
        if not os.path.exists(so):
            return None
        _REF = ctypes.CDLL(so)
        for f in ("ref_sah", "ref_epo", "ref_lcv"):
            if hasattr(_REF, f):
                getattr(_REF, f).restype = ctypes.c_double
    return _REF


_REFCL = None


def refcl():
    """The reference's own OpenCL kernel sources compiled as C++ (oracle/ref_kernels.cpp); None where oracle/_ref
    was never built (no /root/reference and no prebuilt library)."""
    global _REFCL
    if _REFCL is None:
        so = os.path.join(_HERE, "_ref", "libmcpt_refcl.so")
        if not os.path.exists(so):
            subprocess.check_call(["bash", os.path.join(_HERE, "build_ref.sh")])
        if not os.path.exists(so):
            return None
        _REFCL = ctypes.CDLL(so)
    return _REFCL


def refcl_generate_rays(camera, width, height):
    from montecarlopathtracing_b200 import types as T
    rays = np.zeros(width * height, dtype=T.RAY)
    cam = np.ascontiguousarray(camera).copy()
    refcl().refcl_generate_rays(_vp(cam), width, height, _vp(rays))
    return rays


def refcl_intersect(rays, nodes, tris, tmin=0.001, hits=None):
    from montecarlopathtracing_b200 import types as T
    if hits is None:
        hits = np.zeros(len(rays), dtype=T.HIT)
    refcl().refcl_intersect(_vp(rays), _vp(nodes), _vp(tris), _vp(hits), ctypes.c_long(len(rays)), ctypes.c_float(tmin))
    return hits


def refcl_shade(materials, rays, hits, color, seeds, max_depth):
    refcl().refcl_shade(_vp(materials), _vp(rays), _vp(hits), _vp(color), _vp(seeds), ctypes.c_long(len(rays)), max_depth)


def refcl_history(new_colors, histories, counts, width, height, max_attempt):
    refcl().refcl_history(_vp(new_colors), _vp(histories), _vp(counts), width, height, max_attempt)


def refcl_update(camera, width, height, nodes, tris, materials, max_depth, max_attempt, frames, seeds, hist, counts,
                 tmin=0.001):
    """`frames` iterations of OpenCL::update with the reference's kernels; seeds / hist / counts are updated in place.
    Returns the number of live rays that entered intersect."""
    traced = ctypes.c_ulonglong(0)
    cam = np.ascontiguousarray(camera).copy()
    refcl().refcl_update(_vp(cam), width, height, _vp(nodes), _vp(tris), _vp(materials), max_depth, max_attempt, frames,
                         ctypes.c_float(tmin), _vp(seeds), _vp(hist), _vp(counts), ctypes.byref(traced))
    return traced.value


def refcl_display(color, width, height):
    out = np.zeros_like(color)
    refcl().refcl_display(_vp(color), _vp(out), width, height)
    return out


# ---------------------------------------------------------------- BVH quality metrics (SURVEY f-4)
def epo_per_leaf(nodes, tris):
    """Restatement of kernels/EPO.cl: (epo[ntris], area[ntris]) float32, indexed by leaf slot."""
    nodes = np.ascontiguousarray(nodes)
    tris = np.ascontiguousarray(tris)
    epo = np.zeros(len(tris), np.float32)
    area = np.zeros(len(tris), np.float32)
    r = oracle().oracle_epo_per_leaf(_vp(nodes), len(nodes), _vp(tris), len(tris), _vp(epo), _vp(area))
    if r != 0:
        raise RuntimeError("oracle_epo_per_leaf: %d" % r)
    return epo, area


def epo_max_polygon(reset=True):
    """Largest clipped polygon the restatement has seen since the last reset (the reference kernel's buffers hold 8)."""
    return oracle().oracle_epo_max_polygon(int(reset))


def refcl_epo(nodes, tris):
    """kernels/EPO.cl itself (compiled as C++): (epo[ntris], area[ntris])."""
    nodes, tris = np.ascontiguousarray(nodes).copy(), np.ascontiguousarray(tris).copy()
    epo = np.zeros(len(tris), np.float32)
    area = np.zeros(len(tris), np.float32)
    refcl().refcl_epo(_vp(nodes), _vp(tris), _vp(epo), _vp(area), len(tris))
    return epo, area


def epo_total(epo, area):
    return oracle().oracle_epo_total(_vp(np.ascontiguousarray(epo, np.float32)), _vp(np.ascontiguousarray(area, np.float32)), len(epo))


def ref_sah(nodes):
    nodes = np.ascontiguousarray(nodes)
    return ref().ref_sah(_vp(nodes), len(nodes))


def ref_epo(nodes, tris):
    nodes, tris = np.ascontiguousarray(nodes), np.ascontiguousarray(tris)
    return ref().ref_epo(_vp(nodes), len(nodes), _vp(tris), len(tris))


def ref_lcv(nodes, camera, width, height):
    nodes, camera = np.ascontiguousarray(nodes), np.ascontiguousarray(camera)
    return ref().ref_lcv(_vp(nodes), len(nodes), _vp(camera), int(width), int(height))


# ---------------------------------------------------------------- kernel restatements
def generate_rays(camera, width, height):
    from montecarlopathtracing_b200 import types as T
    rays = np.zeros(width * height, dtype=T.RAY)
    oracle().oracle_generate_rays(_vp(camera), width, height, _vp(rays))
    return rays


def intersect(rays, nodes, tris, tmin=0.001, mode=0, count=False, hits=None):
    from montecarlopathtracing_b200 import types as T
    n = len(rays)
    if hits is None:
        hits = np.zeros(n, dtype=T.HIT)
    closest = np.zeros(n, dtype=np.int32)
    counters = np.zeros(2, dtype=np.uint64)
    oracle().oracle_intersect(_vp(rays), ctypes.c_long(n), _vp(nodes), _vp(tris), ctypes.c_float(tmin), _vp(hits),
                              _vp(closest), mode, _vp(counters) if count else None)
    return (hits, closest, counters) if count else (hits, closest)


def shade(materials, rays, hits, color, seeds, max_depth):
    oracle().oracle_shade(_vp(materials), _vp(rays), _vp(hits), _vp(color), _vp(seeds), ctypes.c_long(len(rays)), max_depth)


def history(new_colors, histories, counts, max_attempt):
    oracle().oracle_history(_vp(new_colors), _vp(histories), _vp(counts), ctypes.c_long(len(counts)), max_attempt)


def display(color):
    out = np.zeros_like(color)
    oracle().oracle_display(_vp(color), _vp(out), ctypes.c_long(color.size // 4))
    return out


def init_seeds(n, run_seed):
    s = np.zeros(n, dtype=np.uint32)
    oracle().oracle_init_seeds(_vp(s), ctypes.c_long(n), ctypes.c_uint32(run_seed))
    return s


def render(camera, width, height, nodes, tris, materials, max_depth, max_attempt, frames, seed_mode=0,
           run_seed=1, first_sample=0, tmin=0.001):
    n = width * height
    hist = np.zeros((n, 4), np.float32)
    counts = np.zeros(n, np.int32)
    s = np.zeros((n, 3), np.float64)
    s2 = np.zeros((n, 3), np.float64)
    traced = ctypes.c_uint64(0)
    oracle().oracle_render(_vp(camera), width, height, _vp(nodes), _vp(tris), _vp(materials), max_depth, max_attempt,
                           frames, seed_mode, ctypes.c_uint32(run_seed), ctypes.c_uint32(first_sample),
                           ctypes.c_float(tmin), _vp(hist), _vp(counts), _vp(s), _vp(s2), ctypes.byref(traced))
    return {"hist": hist, "counts": counts, "sum": s, "sumsq": s2, "rays_traced": traced.value}


def set_threads(n=None):
    """Use n OpenMP threads (default: all host cores) regardless of OMP_NUM_THREADS; returns the count in effect."""
    L = oracle()
    L.oracle_set_threads(int(n or os.cpu_count() or 1))
    return L.oracle_max_threads()
