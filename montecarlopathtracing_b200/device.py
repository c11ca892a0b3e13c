"""ctypes binding of libmcpt_cuda.so (include/mcpt.h) — the CUDA hot path.

The classes mirror the reference's stage interface so that tests read like the reference's call
sites: RayGeneration.generateRay (raygeneration.cpp:29-35), SceneBase.intersect / .shade
(scenebuild.cpp:103-145), ColorOut.outputColor (colorout.cpp:40-73) and OpenCLApp.update
(OpenCLApp.cpp:57-82).  There is no CPU fallback: loading fails loudly when the library is missing
and every call raises when no sm_100 device is usable.
"""
import ctypes
import os

import numpy as np

from . import types as T

_PKG_DIR = os.path.dirname(os.path.abspath(__file__))
_LIB = None

TRAVERSAL_REFERENCE, TRAVERSAL_FAST, TRAVERSAL_CANONICAL, TRAVERSAL_WIDE8 = 0, 1, 2, 3
SAMPLING_REFERENCE, SAMPLING_NEE_MIS, SAMPLING_NEE_MIS_RATIO = 0, 1, 2

EXPORTS = [
    "mcpt_last_error", "mcpt_abi_version", "mcpt_create", "mcpt_destroy", "mcpt_stream", "mcpt_synchronize",
    "mcpt_upload_scene", "mcpt_set_camera", "mcpt_generate_rays", "mcpt_reset_color", "mcpt_intersect", "mcpt_shade",
    "mcpt_accumulate", "mcpt_display", "mcpt_frame", "mcpt_render", "mcpt_render_async", "mcpt_set_first_sample",
    "mcpt_reset_accumulation", "mcpt_read_rays", "mcpt_write_rays", "mcpt_read_hits", "mcpt_write_hits",
    "mcpt_read_color", "mcpt_write_color", "mcpt_read_seeds", "mcpt_write_seeds", "mcpt_read_history",
    "mcpt_read_sums", "mcpt_bind_sums", "mcpt_trace", "mcpt_trace_device", "mcpt_trace_count", "mcpt_get_stats",
    "mcpt_reset_stats", "mcpt_get_queue_sorting", "mcpt_set_profiling", "mcpt_get_bounce_time", "mcpt_build_lbvh",
    "mcpt_optimize_treelet", "mcpt_measure_read_bandwidth", "mcpt_bvh_metrics", "mcpt_merge_sums",
    "mcpt_upload_scene_tagged", "mcpt_write_sums", "mcpt_write_history", "mcpt_get_first_sample",
    "mcpt_trace_noalloc", "mcpt_derive_scene", "mcpt_free_blob", "mcpt_upload_scene_derived", "mcpt_phong_accept_table", "mcpt_check_failures", "mcpt_derive_scene_device",
]
TRACE_NOALLOC_MAX_RAYS = 65536        # MCPT_TRACE_NOALLOC_MAX_RAYS of include/mcpt.h


class Params(ctypes.Structure):
    _fields_ = [("width", ctypes.c_int), ("height", ctypes.c_int), ("max_depth", ctypes.c_int),
                ("max_attempt", ctypes.c_int), ("device", ctypes.c_int), ("seed_mode", ctypes.c_int),
                ("run_seed", ctypes.c_uint32), ("samples_in_flight", ctypes.c_int), ("tmin", ctypes.c_float),
                ("traversal", ctypes.c_int), ("sampling", ctypes.c_int), ("reserved", ctypes.c_int * 5)]


class Stats(ctypes.Structure):
    _fields_ = [("rays_traced", ctypes.c_uint64), ("paths", ctypes.c_uint64), ("kernel_launches", ctypes.c_uint64),
                ("nodes_fetched", ctypes.c_uint64), ("tris_fetched", ctypes.c_uint64),
                ("leaves_fetched", ctypes.c_uint64),
                ("slow_rays", ctypes.c_uint64), ("replayed_rays", ctypes.c_uint64)]


def library_path():
    return os.environ.get("MCPT_LIB") or os.path.join(_PKG_DIR, "libmcpt_cuda.so")   # MCPT_LIB: development A/B builds


def lib():
    global _LIB
    if _LIB is None:
        path = library_path()
        if not os.path.exists(path):
            raise RuntimeError("libmcpt_cuda.so is not built - run `python __graft_entry__.py` (there is no CPU fallback)")
        L = ctypes.CDLL(path)
        L.mcpt_last_error.restype = ctypes.c_char_p
        L.mcpt_create.restype = ctypes.c_void_p
        L.mcpt_create.argtypes = [ctypes.POINTER(Params)]
        L.mcpt_stream.restype = ctypes.c_void_p
        L.mcpt_stream.argtypes = [ctypes.c_void_p]
        L.mcpt_destroy.restype = None
        L.mcpt_destroy.argtypes = [ctypes.c_void_p]
        _LIB = L
    return _LIB


def _vp(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(ctypes.c_void_p)
    return ctypes.c_void_p(int(a))       # raw device pointer (e.g. torch.Tensor.data_ptr())


class McptError(RuntimeError):
    pass


class Context:
    """One device context = the reference's OpenCLApp state (OpenCLApp.cpp:17-33) for one scene."""

    def __init__(self, width, height, max_depth, max_attempt, device=0, seed_mode=0, run_seed=1, samples_in_flight=0,
                 tmin=0.001, traversal=TRAVERSAL_FAST, sampling=0):
        self.width, self.height, self.max_depth, self.max_attempt = width, height, max_depth, max_attempt
        self.npix = width * height
        p = Params(width, height, max_depth, max_attempt, device, seed_mode, run_seed, samples_in_flight, tmin, traversal, sampling)
        self._L = lib()
        self._h = self._L.mcpt_create(ctypes.byref(p))
        if not self._h:
            raise McptError("mcpt_create: " + self._L.mcpt_last_error().decode())
        self._h = ctypes.c_void_p(self._h)

    def close(self):
        if getattr(self, "_h", None):
            self._L.mcpt_destroy(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _call(self, name, *args):
        rc = getattr(self._L, name)(self._h, *args)
        if rc != 0:
            raise McptError("%s: %s" % (name, self._L.mcpt_last_error().decode()))

    @property
    def stream(self):
        return self._L.mcpt_stream(self._h)

    def synchronize(self):
        self._call("mcpt_synchronize")

    # ---- OpenCL::init() pieces
    def upload_scene(self, triangles, materials, nodes, geometry_tag=0, derived=None):
        """geometry_tag: the caller's generation counter (non-zero and unchanged = same triangles and nodes as last time).
        derived: a blob of derive_scene() for these very arrays (scene cache): uploaded instead of deriving again."""
        tr = np.ascontiguousarray(triangles, dtype=T.TRIANGLE)
        mt = np.ascontiguousarray(materials, dtype=T.MATERIAL)
        nd = np.ascontiguousarray(nodes, dtype=T.BVH_NODE)
        if derived is not None:
            blob = np.frombuffer(derived, dtype=np.uint8)
            self._call("mcpt_upload_scene_derived", _vp(tr), len(tr), _vp(mt), len(mt), _vp(nd), len(nd), _vp(blob),
                       ctypes.c_size_t(blob.size), ctypes.c_uint64(geometry_tag))
        elif geometry_tag:
            self._call("mcpt_upload_scene_tagged", _vp(tr), len(tr), _vp(mt), len(mt), _vp(nd), len(nd), ctypes.c_uint64(geometry_tag))
        else:
            self._call("mcpt_upload_scene", _vp(tr), len(tr), _vp(mt), len(mt), _vp(nd), len(nd))

    def set_camera(self, camera):
        cam = np.ascontiguousarray(camera, dtype=T.CAMERA)
        self._call("mcpt_set_camera", _vp(cam))

    # ---- stage level (RayGeneration / SceneBase / ColorOut)
    def generate_rays(self):
        self._call("mcpt_generate_rays")

    def reset_color(self):
        self._call("mcpt_reset_color")

    def intersect(self, traversal=TRAVERSAL_FAST):
        self._call("mcpt_intersect", traversal)

    def shade(self):
        self._call("mcpt_shade")

    def accumulate(self):
        self._call("mcpt_accumulate")

    def display(self, from_history=True):
        out = np.zeros((self.height, self.width, 4), np.float32)
        self._call("mcpt_display", int(from_history), _vp(out))
        return out

    def frame(self):
        self._call("mcpt_frame")

    # ---- fused wavefront path
    def render(self, spp):
        self._call("mcpt_render", int(spp))

    def render_async(self, spp):
        self._call("mcpt_render_async", int(spp))

    def set_first_sample(self, first):
        self._call("mcpt_set_first_sample", ctypes.c_uint32(first))

    def reset_accumulation(self):
        self._call("mcpt_reset_accumulation")

    # ---- buffer I/O
    def read_rays(self):
        out = np.zeros(self.npix, T.RAY)
        self._call("mcpt_read_rays", _vp(out))
        return out

    def write_rays(self, rays):
        r = np.ascontiguousarray(rays, dtype=T.RAY)
        assert len(r) == self.npix
        self._call("mcpt_write_rays", _vp(r))

    def merge_sums(self, others):
        """sums += the other contexts' sums (in order), on the device over peer memory (mcpt_merge_sums)."""
        arr = (ctypes.c_void_p * max(1, len(others)))(*[o._h for o in others])
        self._call("mcpt_merge_sums", arr, len(others))

    def read_hits(self):
        hits = np.zeros(self.npix, T.HIT)
        closest = np.zeros(self.npix, np.int32)
        self._call("mcpt_read_hits", _vp(hits), _vp(closest))
        return hits, closest

    def write_hits(self, hits):
        h = np.ascontiguousarray(hits, dtype=T.HIT)
        assert len(h) == self.npix
        self._call("mcpt_write_hits", _vp(h))

    def read_color(self):
        out = np.zeros((self.npix, 4), np.float32)
        self._call("mcpt_read_color", _vp(out))
        return out

    def write_color(self, color):
        c = np.ascontiguousarray(color, dtype=np.float32).reshape(self.npix, 4)
        self._call("mcpt_write_color", _vp(c))

    def read_seeds(self):
        out = np.zeros(self.npix, np.uint32)
        self._call("mcpt_read_seeds", _vp(out))
        return out

    def write_seeds(self, seeds):
        s = np.ascontiguousarray(seeds, dtype=np.uint32)
        assert len(s) == self.npix
        self._call("mcpt_write_seeds", _vp(s))

    def read_history(self):
        hist = np.zeros((self.npix, 4), np.float32)
        counts = np.zeros(self.npix, np.int32)
        self._call("mcpt_read_history", _vp(hist), _vp(counts))
        return hist, counts

    def read_sums(self):
        out = np.zeros((7, self.npix), np.float64)
        self._call("mcpt_read_sums", _vp(out))
        return out

    def write_sums(self, sums):
        s = np.ascontiguousarray(sums, dtype=np.float64).reshape(7, self.npix)
        self._call("mcpt_write_sums", _vp(s))

    def write_history(self, hist, counts):
        h = np.ascontiguousarray(hist, dtype=np.float32).reshape(self.npix, 4)
        c = np.ascontiguousarray(counts, dtype=np.int32).reshape(self.npix)
        self._call("mcpt_write_history", _vp(h), _vp(c))

    def first_sample(self):
        v = ctypes.c_uint32(0)
        self._call("mcpt_get_first_sample", ctypes.byref(v))
        return v.value

    def checkpoint(self):
        """Everything a later context needs to continue this render bit for bit (see restore)."""
        hist, counts = self.read_history()
        return {"hist": hist, "counts": counts, "sums": self.read_sums(), "seeds": self.read_seeds(),
                "first_sample": np.uint32(self.first_sample()), "size": np.array([self.width, self.height])}

    def restore(self, ck, seed_mode):
        assert tuple(int(x) for x in ck["size"]) == (self.width, self.height)
        self.write_history(ck["hist"], ck["counts"])
        self.write_sums(ck["sums"])
        if seed_mode == 0:
            self.write_seeds(ck["seeds"])
        else:
            self.set_first_sample(int(ck["first_sample"]))

    def bind_sums(self, device_ptr):
        self._call("mcpt_bind_sums", _vp(device_ptr))

    # ---- microbench
    def trace(self, rays, traversal=TRAVERSAL_FAST, hits=None, closest=None):
        r = rays if (isinstance(rays, np.ndarray) and rays.dtype == T.RAY and rays.flags.c_contiguous) \
            else np.ascontiguousarray(rays, dtype=T.RAY)
        n = len(r)
        if hits is None:
            hits = np.zeros(n, T.HIT)
        if closest is None:
            closest = np.zeros(n, np.int32)
        self._call("mcpt_trace", _vp(r), ctypes.c_long(n), _vp(hits), _vp(closest), traversal)
        return hits, closest

    def trace_noalloc(self, rays, hits=None, closest=None):
        """mcpt_trace_noalloc: the product walk on host rays with the device scratch in a static array of the library
        (no cudaMalloc inside the call); at most TRACE_NOALLOC_MAX_RAYS rays."""
        r = rays if (isinstance(rays, np.ndarray) and rays.dtype == T.RAY and rays.flags.c_contiguous) \
            else np.ascontiguousarray(rays, dtype=T.RAY)
        n = len(r)
        if hits is None:
            hits = np.zeros(n, T.HIT)
        if closest is None:
            closest = np.zeros(n, np.int32)
        self._call("mcpt_trace_noalloc", _vp(r), ctypes.c_long(n), _vp(hits), _vp(closest))
        return hits, closest

    def trace_device(self, d_rays, n, d_hits, d_closest=None, traversal=TRAVERSAL_FAST):
        self._call("mcpt_trace_device", _vp(d_rays), ctypes.c_long(n), _vp(d_hits), _vp(d_closest), traversal)

    def trace_count(self, d_rays, n, traversal=TRAVERSAL_FAST):
        self._call("mcpt_trace_count", _vp(d_rays), ctypes.c_long(n), traversal)
        return self.stats()

    def stats(self):
        s = Stats()
        self._call("mcpt_get_stats", ctypes.byref(s))
        return {k: getattr(s, k) for k, _ in Stats._fields_}

    def set_profiling(self, count_fetches=False, time_bounces=False):
        self._call("mcpt_set_profiling", int(count_fetches), int(time_bounces))

    def bounce_time(self):
        """(summed device ms of the bounce kernels, launches) since the last call; needs time_bounces."""
        ms, n = ctypes.c_double(0), ctypes.c_uint64(0)
        self._call("mcpt_get_bounce_time", ctypes.byref(ms), ctypes.byref(n))
        return ms.value, n.value

    def reset_stats(self):
        self._call("mcpt_reset_stats")

    def queue_sorting(self):
        """bit 0: path queue sorted, bit 1: shadow-ray queue sorted; -1 while the NEE + MIS modes are still timing both."""
        v = ctypes.c_int(0)
        self._call("mcpt_get_queue_sorting", ctypes.byref(v))
        return v.value


def image_from_sums(sums, width, height):
    """Mean of the counted (non-zero) samples = the reference's image (history.cl), from mergeable sums."""
    n = np.maximum(sums[6], 1.0)
    img = np.zeros((height, width, 4), np.float32)
    img[..., :3] = (sums[:3] / n).T.reshape(height, width, 3)
    return img


def derive_scene(triangles, materials, nodes, device=None, return_ms=False):
    """The traversal layout mcpt_upload_scene derives (leaf records, two-box nodes, the binned-SAH tree of the fast
    walk) as one relocatable blob (bytes) - what the scene cache stores next to the arrays.  device=None: all on the
    host, needs no GPU; device=k: the tree is built on that GPU (mcpt_derive_scene_device) - the same bytes."""
    tr = np.ascontiguousarray(triangles, dtype=T.TRIANGLE)
    mt = np.ascontiguousarray(materials, dtype=T.MATERIAL)
    nd = np.ascontiguousarray(nodes, dtype=T.BVH_NODE)
    blob, n = ctypes.c_void_p(), ctypes.c_size_t()
    L = lib()
    L.mcpt_free_blob.restype = None
    L.mcpt_free_blob.argtypes = [ctypes.c_void_p]
    ms = ctypes.c_float(0.0)
    if device is None:
        rc = L.mcpt_derive_scene(_vp(tr), len(tr), _vp(mt), len(mt), _vp(nd), len(nd), ctypes.byref(blob), ctypes.byref(n))
    else:
        rc = L.mcpt_derive_scene_device(int(device), _vp(tr), len(tr), _vp(mt), len(mt), _vp(nd), len(nd), ctypes.byref(blob),
                                        ctypes.byref(n), ctypes.byref(ms))
    if rc != 0:
        raise McptError("mcpt_derive_scene: %s" % L.mcpt_last_error().decode())
    try:
        data = ctypes.string_at(blob, n.value)
        return (data, ms.value) if return_ms else data
    finally:
        L.mcpt_free_blob(blob)


def check_failures(device=0):
    """Checked build only (MCPT_LIB=build/ab/libmcpt_checked.so): (failures, first code, first value, first bound)."""
    out = (ctypes.c_uint64 * 4)()
    if lib().mcpt_check_failures(int(device), out) != 0:
        raise McptError("mcpt_check_failures: %s" % lib().mcpt_last_error().decode())
    return tuple(int(x) for x in out)


def phong_accept_table():
    """Acceptance probability of the glossy lobe's rejection loop against angle(mirror direction, normal) (NEE + MIS mode)."""
    out = np.zeros(1025, np.float32)
    if lib().mcpt_phong_accept_table(_vp(out), len(out)) != 0:
        raise McptError("mcpt_phong_accept_table: %s" % lib().mcpt_last_error().decode())
    return out


def build_lbvh(triangles, device=0, return_ms=False):
    """HLBVH node array built on the GPU (mcpt_build_lbvh): byte-identical to host.build_bvh(triangles)."""
    tris = np.ascontiguousarray(triangles, dtype=T.TRIANGLE)
    nodes = np.zeros(max(1, 2 * len(tris) - 1), dtype=T.BVH_NODE)
    ms = ctypes.c_float(0.0)
    r = lib().mcpt_build_lbvh(int(device), _vp(tris), len(tris), _vp(nodes), ctypes.byref(ms))
    if r < 0:
        raise McptError("mcpt_build_lbvh: %s" % lib().mcpt_last_error().decode())
    assert r == len(nodes)
    return (nodes, ms.value) if return_ms else nodes


class BvhQuality(ctypes.Structure):
    _fields_ = [("sah", ctypes.c_double), ("epo", ctypes.c_double), ("lcv", ctypes.c_double),
                ("sah_exact", ctypes.c_double), ("epo_area", ctypes.c_double), ("triangle_area", ctypes.c_double),
                ("lcv_mean_leaves", ctypes.c_double), ("lcv_rays", ctypes.c_uint64),
                ("ms_sah", ctypes.c_float), ("ms_epo", ctypes.c_float), ("ms_lcv", ctypes.c_float), ("pad_", ctypes.c_float)]


def bvh_metrics(nodes, triangles=None, camera=None, width=0, height=0, device=0, per_leaf=False):
    """SAH / EPO / LCV of a node array on the GPU (mcpt_bvh_metrics); dict of the record's fields
    (+ 'epo_per_leaf', 'area_per_leaf' float32 arrays with per_leaf=True)."""
    nd = np.ascontiguousarray(nodes, dtype=T.BVH_NODE)
    tr = np.ascontiguousarray(triangles, dtype=T.TRIANGLE) if triangles is not None else None
    cam = np.ascontiguousarray(camera) if camera is not None else None
    q = BvhQuality()
    epo = np.zeros(len(tr), np.float32) if (per_leaf and tr is not None) else None
    area = np.zeros(len(tr), np.float32) if (per_leaf and tr is not None) else None
    r = lib().mcpt_bvh_metrics(int(device), _vp(nd), len(nd), _vp(tr) if tr is not None else None,
                               len(tr) if tr is not None else 0, _vp(cam) if cam is not None else None,
                               int(width), int(height), ctypes.byref(q),
                               _vp(epo) if epo is not None else None, _vp(area) if area is not None else None)
    if r < 0:
        raise McptError("mcpt_bvh_metrics: %s" % lib().mcpt_last_error().decode())
    out = {k: getattr(q, k) for k, _ in BvhQuality._fields_ if k != "pad_"}
    if per_leaf:
        out["epo_per_leaf"], out["area_per_leaf"] = epo, area
    return out


def measure_read_bandwidth(nbytes, iters=20, device=0):
    """GB/s of a streaming read over a scratch buffer of nbytes (32 MiB: L2 ceiling; GiBs: HBM ceiling)."""
    g = ctypes.c_float(0.0)
    f = lib().mcpt_measure_read_bandwidth
    f.argtypes = [ctypes.c_int, ctypes.c_size_t, ctypes.c_int, ctypes.POINTER(ctypes.c_float)]
    if f(int(device), int(nbytes), int(iters), ctypes.byref(g)) < 0:
        raise McptError("mcpt_measure_read_bandwidth: %s" % lib().mcpt_last_error().decode())
    return g.value


def optimize_treelet(nodes, device=0, return_ms=False):
    """Treelet-restructured node array from the GPU (mcpt_optimize_treelet): byte-identical to host.treelet(nodes)."""
    src = np.ascontiguousarray(nodes, dtype=T.BVH_NODE)
    out = np.zeros_like(src)
    ms = ctypes.c_float(0.0)
    r = lib().mcpt_optimize_treelet(int(device), _vp(src), len(src), _vp(out), ctypes.byref(ms))
    if r < 0:
        raise McptError("mcpt_optimize_treelet: %s" % lib().mcpt_last_error().decode())
    assert r == len(out)
    return (out, ms.value) if return_ms else out
