"""g.p.u-pathtracer_b200 -- Python host mirror of the C-ABI in include/gpt.h.

Thin ctypes layer over ``libgpt_b200.so`` (hand-written sm_100a CUDA, built in-tree by
``g.p.u-pathtracer_b200/Makefile``).  Names follow the reference: ``kernelInfo`` / ``CamInfo`` /
``BVHData`` / ``Sphere`` (GpuPathTracer/CpuStructs.hpp:19-72, CommomStructs.hpp:11-48) and
``BasicScene.launchKernel`` (GpuPathTracer/tracer.cu:405-415, BasicScene.cpp:368-477).

PyTorch is used only as plumbing (device buffers, streams, torch.distributed); all rendering
happens inside the shared library.  There is no CPU or PyTorch fallback: a missing library or a
missing GPU raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgpt_b200.so")

# enums: CommomStructs.hpp:11-12 (order is ABI)
TRI, SPHERE, BOX, NONE = 0, 1, 2, 3
DIFF, METAL, SPEC, REFR = 0, 1, 2, 3

FLAG_NO_JITTER = 1 << 0
FLAG_METAL_FIXED = 1 << 1
FLAG_MEGAKERNEL = 1 << 2
FLAG_COUNT_BOUNCES = 1 << 3
FLAG_TIME_KERNELS = 1 << 4
FLAG_WIDE_BVH = 1 << 5        # always traverse the 4-wide quantised tree
FLAG_BINARY_BVH = 1 << 6      # always traverse the binary tree as uploaded
FLAG_WAVEFRONT = 1 << 7       # always the wavefront pipeline (default: megakernel for calls under GPT_SMALL_FRAME_PATHS paths)
FLAG_BUILD_LBVH = 1 << 8      # build_bvh: Morton radix tree instead of PLOC clustering
FLAG_FAST = 1 << 10           # opt-in fast shading arithmetic (statistical parity, not bit-exact)
FLAG_NO_TRIAL = 1 << 9        # no timed per-scene trial: render calls never synchronise (graph capture, strict asynchrony)
FLAG_KEEP_TREE = 1 << 11      # traverse the tree exactly as uploaded: no second tree from the library's own builder
STATIC_MAX_PIXELS = 256 * 1024   # gpt_render_static (include/gpt.h GPT_STATIC_*): frame pixels, paths per wave, depth
STATIC_MAX_PATHS = 256 * 1024
STATIC_MAX_DEPTH = 64
K_GENERATE, K_EXTEND, K_SHADE, K_COLLECT, K_ACCUMULATE, K_RESOLVE, K_OTHER = range(7)


class GptError(RuntimeError):
    pass


class Sphere(C.Structure):  # CommomStructs.hpp:18-39
    _fields_ = [("posRad", C.c_float * 4), ("emi", C.c_float * 3), ("col", C.c_float * 3), ("mat", C.c_int32)]


class Box(C.Structure):  # CommomStructs.hpp:42-48
    _fields_ = [("mn", C.c_float * 3), ("mx", C.c_float * 3), ("emi", C.c_float * 3), ("col", C.c_float * 3), ("mat", C.c_int32)]


class CamInfo(C.Structure):  # CpuStructs.hpp:19-29
    _fields_ = [("front", C.c_float * 3), ("right", C.c_float * 3), ("up", C.c_float * 3), ("pos", C.c_float * 3),
                ("dist", C.c_float), ("pitch", C.c_float), ("yaw", C.c_float), ("aspect", C.c_float), ("fov", C.c_float),
                ("dirty", C.c_uint8), ("_pad0", C.c_uint8 * 3), ("bias", C.c_float * 2),
                ("enabled", C.c_uint8), ("_pad1", C.c_uint8 * 3)]


class BVHData(C.Structure):  # CpuStructs.hpp:31-42
    _fields_ = [("dev_triNode", C.c_void_p), ("triNodeSize", C.c_size_t),
                ("dev_triIndicesPtr", C.c_void_p), ("triIndicesSize", C.c_size_t),
                ("dev_triPtr", C.c_void_p), ("triSize", C.c_size_t)]


class kernelInfo(C.Structure):  # CpuStructs.hpp:45-72 (296 bytes, bit-compatible)
    _fields_ = [("blockSize", C.c_uint32 * 3), ("_pad0", C.c_uint32),
                ("hash", C.c_uint64), ("dev_drawRes", C.c_void_p), ("accumBuffer", C.c_void_p),
                ("width", C.c_int32), ("height", C.c_int32), ("cam", CamInfo), ("_pad1", C.c_uint32),
                ("triangleTex", C.c_void_p), ("numVerts", C.c_size_t),
                ("sphereTex", C.c_void_p), ("numSpheres", C.c_size_t),
                ("cullBackFaces", C.c_uint8), ("_pad2", C.c_uint8 * 3), ("depth", C.c_uint32),
                ("constantPdf", C.c_uint64), ("bvhData", BVHData),
                ("time_elapsed", C.c_float), ("air_ref_index", C.c_float), ("glass_ref_index", C.c_float),
                ("triCurrentMat", C.c_int32), ("clearBuffer", C.c_uint8), ("_pad3", C.c_uint8 * 3),
                ("bkColor", C.c_float * 3), ("phongExpo", C.c_float), ("emi", C.c_float * 3), ("col", C.c_float * 3),
                ("_pad4", C.c_uint32)]


class Stats(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("rays_by_bounce", C.c_uint64 * 16), ("paths", C.c_uint64), ("kernel_launches", C.c_uint64)]


class TriMaterial(C.Structure):  # gpt_tri_material (N4)
    _fields_ = [("col", C.c_float * 3), ("emi", C.c_float * 3), ("mat", C.c_int32)]


class BuiltBVH(C.Structure):  # gpt_built_bvh
    _fields_ = [("bvh", BVHData), ("n_inner_nodes", C.c_uint32), ("n_leaves", C.c_uint32), ("n_triangles", C.c_uint32),
                ("depth", C.c_uint32), ("build_ms", C.c_float)]


class Peaks(C.Structure):  # gpt_peaks
    _fields_ = [("gather_l1_gbs", C.c_double), ("gather_l2_gbs", C.c_double), ("gather_hbm_gbs", C.c_double), ("warp_inst_per_s", C.c_double),
                ("stream_l1_gbs", C.c_double), ("stream_l2_gbs", C.c_double)]


class KernelTimes(C.Structure):
    _fields_ = [("ms", C.c_double * 8), ("launches", C.c_uint64 * 8)]


assert C.sizeof(Sphere) == 44 and C.sizeof(Box) == 52 and C.sizeof(CamInfo) == 84
assert C.sizeof(BVHData) == 48 and C.sizeof(kernelInfo) == 296

# every symbol include/gpt.h declares (tests check the library exports each one)
EXPORTS = ["gpt_create", "gpt_destroy", "gpt_set_flags", "gpt_last_error_string", "gpt_version", "gpt_sync",
           "gpt_upload_scene", "gpt_launch", "gpt_invalidate_scene", "gpt_render", "gpt_render_static", "gpt_resolve", "gpt_render_frame_host",
           "gpt_trace_primary", "gpt_trace_rays", "gpt_get_stats", "gpt_reset_stats", "gpt_get_kernel_times", "gpt_measure_peaks", "gpt_wide_info", "gpt_reorder_info", "gpt_trial_info", "gpt_rebuild_info", "gpt_tree_info", "gpt_megakernel_tree_info",
           "gpt_build_bvh", "gpt_free_bvh", "gpt_copy_to_host",
           "gpt_frame_alloc", "gpt_frame_open_peers", "gpt_frame_ptr", "gpt_frame_clear", "gpt_reduce_resolve",
           "gpt_reduce_resolve_ptrs", "gpt_frame_free",
           "gpt_set_triangle_materials", "gpt_load_obj", "gpt_load_obj_materials", "gpt_free_host", "gpt_write_png", "gpt_write_pfm", "gpt_write_exr", "gpt_save_accum", "gpt_load_accum",
           "gpt_hash64", "gpt_default_spheres", "gpt_default_kernel_info", "gpt_set_camera"]

_lib = None


def load_library():
    """Load libgpt_b200.so (built by the in-tree Makefile). Raises if it is missing: no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise GptError("%s not found: build it with `make -C %s` (or __graft_entry__.build()); "
                       "there is no CPU/PyTorch fallback" % (LIB_PATH, _HERE))
    L = C.CDLL(LIB_PATH)
    L.gpt_create.argtypes = [C.POINTER(C.c_void_p), C.c_int]
    L.gpt_destroy.argtypes = [C.c_void_p]
    L.gpt_destroy.restype = None
    L.gpt_set_flags.argtypes = [C.c_void_p, C.c_uint32]
    L.gpt_last_error_string.argtypes = [C.c_void_p]
    L.gpt_last_error_string.restype = C.c_char_p
    L.gpt_version.restype = C.c_char_p
    L.gpt_sync.argtypes = [C.c_void_p, C.c_void_p]
    L.gpt_upload_scene.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t,
                                   C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_int]
    L.gpt_launch.argtypes = [C.c_void_p, C.POINTER(kernelInfo), C.c_void_p]
    L.gpt_invalidate_scene.argtypes = [C.c_void_p]
    L.gpt_render.argtypes = [C.c_void_p, C.POINTER(kernelInfo), C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_void_p]
    L.gpt_render_static.argtypes = L.gpt_render.argtypes
    L.gpt_resolve.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    L.gpt_render_frame_host.argtypes = [C.c_void_p, C.POINTER(kernelInfo), C.c_uint64, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p]
    L.gpt_trace_primary.argtypes = [C.c_void_p, C.POINTER(kernelInfo), C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.gpt_trace_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.gpt_get_stats.argtypes = [C.c_void_p, C.POINTER(Stats)]
    L.gpt_reset_stats.argtypes = [C.c_void_p]
    L.gpt_get_kernel_times.argtypes = [C.c_void_p, C.POINTER(KernelTimes)]
    L.gpt_measure_peaks.argtypes = [C.c_void_p, C.POINTER(Peaks)]
    L.gpt_wide_info.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.gpt_reorder_info.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.gpt_trial_info.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_double)]
    L.gpt_tree_info.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_double), C.POINTER(C.c_int)]
    L.gpt_megakernel_tree_info.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_double)]
    L.gpt_rebuild_info.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_double), C.POINTER(C.c_float)]
    L.gpt_build_bvh.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_int, C.c_uint32, C.POINTER(BuiltBVH)]
    L.gpt_free_bvh.argtypes = [C.c_void_p, C.POINTER(BuiltBVH)]
    L.gpt_copy_to_host.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
    L.gpt_hash64.argtypes = [C.c_uint64]
    L.gpt_hash64.restype = C.c_uint64
    L.gpt_default_spheres.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
    L.gpt_default_kernel_info.argtypes = [C.POINTER(kernelInfo), C.c_int, C.c_int]
    L.gpt_default_kernel_info.restype = None
    L.gpt_set_camera.argtypes = [C.POINTER(CamInfo), C.c_void_p, C.c_float, C.c_float]
    L.gpt_set_camera.restype = None
    L.gpt_frame_alloc.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
    L.gpt_frame_open_peers.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
    L.gpt_frame_ptr.argtypes = [C.c_void_p, C.c_int]
    L.gpt_frame_ptr.restype = C.c_void_p
    L.gpt_frame_clear.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    L.gpt_reduce_resolve.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.gpt_reduce_resolve_ptrs.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.gpt_frame_free.argtypes = [C.c_void_p]
    L.gpt_set_triangle_materials.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
    L.gpt_load_obj.argtypes = [C.c_char_p, C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.c_size_t),
                               C.POINTER(C.POINTER(C.c_int32)), C.POINTER(C.c_size_t)]
    L.gpt_load_obj_materials.argtypes = [C.c_char_p, C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.c_size_t),
                                         C.POINTER(C.POINTER(C.c_int32)), C.POINTER(C.c_size_t),
                                         C.POINTER(C.POINTER(C.c_int32)), C.POINTER(C.POINTER(TriMaterial)), C.POINTER(C.c_size_t)]
    L.gpt_free_host.argtypes = [C.c_void_p]
    L.gpt_free_host.restype = None
    L.gpt_write_png.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_int]
    L.gpt_write_pfm.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_int]
    L.gpt_write_exr.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_int]
    L.gpt_save_accum.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_uint64]
    L.gpt_load_accum.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    _lib = L
    return L


# ---------------------------------------------------------------- host-only helpers
def hash64(key):
    """uf::hash, utilfun.cpp:380-389 (per-frame seed)."""
    return int(load_library().gpt_hash64(C.c_uint64(key & 0xFFFFFFFFFFFFFFFF)))


def default_spheres():
    """The 8 hard-coded spheres of BasicScene.cpp:177-218 as a ctypes array."""
    n = C.c_size_t(0)
    L = load_library()
    L.gpt_default_spheres(None, 0, C.byref(n))
    arr = (Sphere * n.value)()
    rc = L.gpt_default_spheres(arr, n.value, C.byref(n))
    if rc != 0:
        raise GptError("gpt_default_spheres rc=%d" % rc)
    return arr


def default_kernel_info(width, height):
    """kernelInfo filled like BasicScene's constructor (BasicScene.cpp:220-259)."""
    info = kernelInfo()
    load_library().gpt_default_kernel_info(C.byref(info), int(width), int(height))
    return info


def set_camera(info, pos=(0.0, 0.0, 0.0), yaw=0.0, pitch=0.0):
    p = (C.c_float * 3)(*pos)
    load_library().gpt_set_camera(C.byref(info.cam), p, C.c_float(yaw), C.c_float(pitch))


def spheres_from_list(items):
    """items: iterable of (posRad[4], emi[3], col[3], mat)."""
    items = list(items)
    arr = (Sphere * len(items))()
    for s, (pr, emi, col, mat) in zip(arr, items):
        s.posRad[:] = pr
        s.emi[:] = emi
        s.col[:] = col
        s.mat = int(mat)
    return arr


def load_obj(path):
    """uf::loadIndexedTris (utilfun.cpp:454-524) -> (verts (n,3) float32, faces (m,3) int32)."""
    L = load_library()
    v, t = C.POINTER(C.c_float)(), C.POINTER(C.c_int32)()
    nv, nt = C.c_size_t(0), C.c_size_t(0)
    rc = L.gpt_load_obj(os.fsencode(path), C.byref(v), C.byref(nv), C.byref(t), C.byref(nt))
    if rc != 0:
        raise GptError("gpt_load_obj(%s) failed (%d)" % (path, rc))
    try:
        verts = np.ctypeslib.as_array(v, shape=(nv.value * 3,)).copy().reshape(-1, 3) if nv.value else np.zeros((0, 3), np.float32)
        faces = np.ctypeslib.as_array(t, shape=(nt.value * 3,)).copy().reshape(-1, 3) if nt.value else np.zeros((0, 3), np.int32)
    finally:
        L.gpt_free_host(v)
        L.gpt_free_host(t)
    return verts, faces


def load_obj_materials(path):
    """OBJ + .mtl -> (verts, faces, face_material int32 (m,), materials (TriMaterial * n))."""
    L = load_library()
    v, t, fm, mt = C.POINTER(C.c_float)(), C.POINTER(C.c_int32)(), C.POINTER(C.c_int32)(), C.POINTER(TriMaterial)()
    nv, nt, nm = C.c_size_t(0), C.c_size_t(0), C.c_size_t(0)
    rc = L.gpt_load_obj_materials(os.fsencode(path), C.byref(v), C.byref(nv), C.byref(t), C.byref(nt), C.byref(fm), C.byref(mt), C.byref(nm))
    if rc != 0:
        raise GptError("gpt_load_obj_materials(%s) failed (%d)" % (path, rc))
    try:
        verts = np.ctypeslib.as_array(v, shape=(nv.value * 3,)).copy().reshape(-1, 3) if nv.value else np.zeros((0, 3), np.float32)
        faces = np.ctypeslib.as_array(t, shape=(nt.value * 3,)).copy().reshape(-1, 3) if nt.value else np.zeros((0, 3), np.int32)
        face_mat = np.ctypeslib.as_array(fm, shape=(nt.value,)).copy() if nt.value else np.zeros((0,), np.int32)
        mats = (TriMaterial * nm.value)()
        if nm.value:
            C.memmove(mats, mt, nm.value * C.sizeof(TriMaterial))
    finally:
        for q in (v, t, fm, mt):
            L.gpt_free_host(q)
    return verts, faces, face_mat, mats


def write_png(path, rgba8, width, height):
    """rgba8: host uint32/int32 array of W*H 0x00BBGGRR pixels, row 0 = bottom (the dev_drawRes layout)."""
    a = np.ascontiguousarray(rgba8).view(np.uint32)
    rc = load_library().gpt_write_png(os.fsencode(path), a.ctypes.data, int(width), int(height))
    if rc != 0:
        raise GptError("gpt_write_png failed (%d)" % rc)


def write_pfm(path, rgb, width, height):
    a = np.ascontiguousarray(rgb, dtype=np.float32)
    rc = load_library().gpt_write_pfm(os.fsencode(path), a.ctypes.data, int(width), int(height))
    if rc != 0:
        raise GptError("gpt_write_pfm failed (%d)" % rc)


def write_exr(path, rgb, width, height):
    """rgb: host float32 W*H*3, row 0 = bottom (gpt_resolve's out_rgb_f32) -> uncompressed OpenEXR, top row first."""
    a = np.ascontiguousarray(rgb, dtype=np.float32)
    rc = load_library().gpt_write_exr(os.fsencode(path), a.ctypes.data, int(width), int(height))
    if rc != 0:
        raise GptError("gpt_write_exr failed (%d)" % rc)


def save_accum(path, accum_f4, samples_done, seed):
    a = np.ascontiguousarray(accum_f4, dtype=np.float32)
    h, w = a.shape[0], a.shape[1]
    rc = load_library().gpt_save_accum(os.fsencode(path), a.ctypes.data, w, h, C.c_uint64(samples_done), C.c_uint64(seed))
    if rc != 0:
        raise GptError("gpt_save_accum failed (%d)" % rc)


def load_accum(path, width, height):
    """-> (accum (H,W,4) float32, samples_done, seed)"""
    a = np.empty((height, width, 4), np.float32)
    n, s = C.c_uint64(0), C.c_uint64(0)
    rc = load_library().gpt_load_accum(os.fsencode(path), a.ctypes.data, int(width), int(height), C.byref(n), C.byref(s))
    if rc != 0:
        raise GptError("gpt_load_accum failed (%d)" % rc)
    return a, n.value, s.value


def _ptr(x):
    """numpy array / torch tensor / ctypes array / int / None -> raw address."""
    if x is None:
        return None
    if isinstance(x, int):
        return x
    if isinstance(x, np.ndarray):
        return x.ctypes.data
    if hasattr(x, "data_ptr"):
        return x.data_ptr()
    return C.addressof(x)


class Context:
    """One render context on one GPU (gpt_create .. gpt_destroy)."""

    def __init__(self, device=0, flags=0):
        self._L = load_library()
        h = C.c_void_p()
        rc = self._L.gpt_create(C.byref(h), int(device))
        if rc != 0:
            raise GptError("gpt_create failed (%d): %s" % (rc, self._L.gpt_last_error_string(None).decode()))
        self._h = h
        self.device = int(device)
        self._keep = []
        if flags:
            self.set_flags(flags)

    def close(self):
        if getattr(self, "_h", None):
            self._L.gpt_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc, what):
        if rc != 0:
            raise GptError("%s failed (%d): %s" % (what, rc, self._L.gpt_last_error_string(self._h).decode()))

    def set_flags(self, flags):
        self._ck(self._L.gpt_set_flags(self._h, C.c_uint32(flags)), "gpt_set_flags")

    def sync(self, stream=None):
        self._ck(self._L.gpt_sync(self._h, _ptr(stream)), "gpt_sync")

    def upload_scene(self, nodes=None, tris=None, indices=None, spheres=None, boxes=None, on_device=False):
        """nodes/tris: (n,4) float32, indices: int32 -- the CudaBVH arrays; spheres/boxes: ctypes arrays."""
        def cnt(a, div=1):
            if a is None:
                return 0
            n = a.numel() if hasattr(a, "numel") else a.size
            return int(n) // div
        ns = len(spheres) if spheres is not None else 0
        nb = len(boxes) if boxes is not None else 0
        self._ck(self._L.gpt_upload_scene(self._h, _ptr(nodes), cnt(nodes, 4), _ptr(tris), cnt(tris, 4), _ptr(indices), cnt(indices),
                                          _ptr(spheres) if ns else None, ns, _ptr(boxes) if nb else None, nb, 1 if on_device else 0),
                 "gpt_upload_scene")

    def launch(self, info, stream=None):
        """Drop-in for BasicScene::launchKernel (tracer.cu:405-415): one legacy 1-spp pass."""
        self._ck(self._L.gpt_launch(self._h, C.byref(info), _ptr(stream)), "gpt_launch")

    def invalidate_scene(self):
        """The caller rewrote kernelInfo.bvhData's arrays in place: the next launch() re-packs them (gpt.h, pointer-identity rule)."""
        self._ck(self._L.gpt_invalidate_scene(self._h), "gpt_invalidate_scene")

    def render(self, info, accum_f4, seed, sample_begin, sample_count, stream=None):
        self._ck(self._L.gpt_render(self._h, C.byref(info), _ptr(accum_f4), C.c_uint64(seed), C.c_uint64(sample_begin),
                                    C.c_uint32(sample_count), _ptr(stream)), "gpt_render")

    def render_static(self, info, accum_f4, seed, sample_begin, sample_count, stream=None):
        """gpt_render on the library's fixed device scratch: allocates nothing; frames of <= STATIC_MAX_PIXELS, depth <= STATIC_MAX_DEPTH"""
        self._ck(self._L.gpt_render_static(self._h, C.byref(info), _ptr(accum_f4), C.c_uint64(seed), C.c_uint64(sample_begin),
                                           C.c_uint32(sample_count), _ptr(stream)), "gpt_render_static")

    def resolve(self, accum_f4, width, height, out_rgba8=None, out_rgb_f32=None, stream=None):
        self._ck(self._L.gpt_resolve(self._h, _ptr(accum_f4), int(width), int(height), _ptr(out_rgba8), _ptr(out_rgb_f32), _ptr(stream)),
                 "gpt_resolve")

    def render_frame_host(self, info, seed, sample_begin, spp, host_rgba8=None, host_accum_f4=None):
        self._ck(self._L.gpt_render_frame_host(self._h, C.byref(info), C.c_uint64(seed), C.c_uint64(sample_begin), C.c_uint32(spp),
                                               _ptr(host_rgba8), _ptr(host_accum_f4)), "gpt_render_frame_host")

    def trace_primary(self, info, seed, sample, geom=None, tri_idx=None, prim_idx=None, t=None, stream=None):
        self._ck(self._L.gpt_trace_primary(self._h, C.byref(info), C.c_uint64(seed), C.c_uint64(sample),
                                           _ptr(geom), _ptr(tri_idx), _ptr(prim_idx), _ptr(t), _ptr(stream)), "gpt_trace_primary")

    def trace_rays(self, o, d, n, cull, tri_idx=None, tri_addr=None, t=None, stream=None):
        self._ck(self._L.gpt_trace_rays(self._h, _ptr(o), _ptr(d), int(n), int(bool(cull)), _ptr(tri_idx), _ptr(tri_addr), _ptr(t), _ptr(stream)),
                 "gpt_trace_rays")

    def stats(self):
        s = Stats()
        self._ck(self._L.gpt_get_stats(self._h, C.byref(s)), "gpt_get_stats")
        return s

    def build_bvh(self, verts, faces, on_device=False, max_leaf=8):
        """Device-side BVH build (gpt_build_bvh; PLOC topology, Morton radix tree with FLAG_BUILD_LBVH): verts (n,3) float32, faces (m,3) int32 -> BuiltBVH whose .bvh holds
        DEVICE arrays in CudaBVH's compact layout.  Free with free_bvh()."""
        def cnt(a):
            return int(a.numel() if hasattr(a, "numel") else a.size) // 3
        b = BuiltBVH()
        self._ck(self._L.gpt_build_bvh(self._h, _ptr(verts), cnt(verts), _ptr(faces), cnt(faces), 1 if on_device else 0,
                                       C.c_uint32(max_leaf), C.byref(b)), "gpt_build_bvh")
        return b

    def free_bvh(self, b):
        self._ck(self._L.gpt_free_bvh(self._h, C.byref(b)), "gpt_free_bvh")

    def built_to_host(self, b):
        """numpy copies of a BuiltBVH's arrays: dict(nodes (n,4) f32, tris (n,4) f32, indices i32)."""
        out = {}
        for key, ptr, n, dt, w in (("nodes", b.bvh.dev_triNode, b.bvh.triNodeSize, np.float32, 4),
                                   ("tris", b.bvh.dev_triPtr, b.bvh.triSize, np.float32, 4),
                                   ("indices", b.bvh.dev_triIndicesPtr, b.bvh.triIndicesSize, np.int32, 1)):
            a = np.empty((int(n), w) if w > 1 else (int(n),), dtype=dt)
            self._ck(self._L.gpt_copy_to_host(self._h, a.ctypes.data, ptr, a.nbytes), "gpt_copy_to_host")
            out[key] = a
        return out

    def upload_built(self, b, spheres=None, boxes=None):
        """gpt_upload_scene(on_device=1) with a BuiltBVH's arrays."""
        ns = len(spheres) if spheres is not None else 0
        nb = len(boxes) if boxes is not None else 0
        self._ck(self._L.gpt_upload_scene(self._h, b.bvh.dev_triNode, b.bvh.triNodeSize, b.bvh.dev_triPtr, b.bvh.triSize,
                                          b.bvh.dev_triIndicesPtr, b.bvh.triIndicesSize, _ptr(spheres) if ns else None, ns,
                                          _ptr(boxes) if nb else None, nb, 1), "gpt_upload_scene")

    # ---- multi-GPU peer frames (gpt_frame_* / gpt_reduce_resolve) ----
    def frame_alloc(self, width, height, n_buffers=2):
        """-> 64-byte IPC handle (bytes) of this rank's frame allocation"""
        h = (C.c_ubyte * 64)()
        self._ck(self._L.gpt_frame_alloc(self._h, int(width), int(height), int(n_buffers), h), "gpt_frame_alloc")
        return bytes(h)

    def frame_open_peers(self, handles, rank):
        """handles: list of 64-byte handles, one per rank (this rank's own entry is ignored)"""
        blob = (C.c_ubyte * (64 * len(handles)))(*b"".join(handles))
        self._ck(self._L.gpt_frame_open_peers(self._h, blob, len(handles), int(rank)), "gpt_frame_open_peers")

    def frame_ptr(self, index):
        p = self._L.gpt_frame_ptr(self._h, int(index))
        if not p:
            raise GptError("gpt_frame_ptr: no such frame buffer")
        return p

    def frame_clear(self, index, stream=None):
        self._ck(self._L.gpt_frame_clear(self._h, int(index), _ptr(stream)), "gpt_frame_clear")

    def reduce_resolve(self, index, total_accum_f4=None, out_rgba8=None, out_rgb_f32=None, stream=None):
        self._ck(self._L.gpt_reduce_resolve(self._h, int(index), _ptr(total_accum_f4), _ptr(out_rgba8), _ptr(out_rgb_f32), _ptr(stream)),
                 "gpt_reduce_resolve")

    def reduce_resolve_ptrs(self, frames, width, height, total_accum_f4=None, out_rgba8=None, out_rgb_f32=None, stream=None):
        arr = (C.c_void_p * len(frames))(*[_ptr(f) for f in frames])
        self._ck(self._L.gpt_reduce_resolve_ptrs(self._h, arr, len(frames), int(width), int(height), _ptr(total_accum_f4), _ptr(out_rgba8),
                                                 _ptr(out_rgb_f32), _ptr(stream)), "gpt_reduce_resolve_ptrs")

    def frame_free(self):
        self._ck(self._L.gpt_frame_free(self._h), "gpt_frame_free")

    def set_triangle_materials(self, materials, face_material):
        """materials: (TriMaterial * n) array; face_material: int32 array indexed by OBJ face index (None clears)."""
        if face_material is None:
            self._ck(self._L.gpt_set_triangle_materials(self._h, None, 0, None, 0), "gpt_set_triangle_materials")
            return
        fm = np.ascontiguousarray(face_material, dtype=np.int32)
        self._ck(self._L.gpt_set_triangle_materials(self._h, C.addressof(materials), len(materials), fm.ctypes.data, fm.size),
                 "gpt_set_triangle_materials")

    def wide_info(self):
        """(wide nodes, wide levels) of the 4-wide tree in use; (0, 0) = the binary tree is traversed."""
        n, lv = C.c_int(0), C.c_int(0)
        self._ck(self._L.gpt_wide_info(self._h, C.byref(n), C.byref(lv)), "gpt_wide_info")
        return n.value, lv.value

    def reorder_info(self):
        """(state, ns/ray in path order, ns/ray re-ordered): state 1 = rays are re-ordered before secondary extends, 0 = not,
        -1 = the per-scene trial is still running."""
        st, a, b = C.c_int(0), C.c_double(0.0), C.c_double(0.0)
        self._ck(self._L.gpt_reorder_info(self._h, C.byref(st), C.byref(a), C.byref(b)), "gpt_reorder_info")
        return st.value, a.value, b.value

    def trial_info(self):
        """(pending, [ns/ray binary tree, ns/ray wide tree, ns/ray chosen tree re-ordered]) of the per-scene trial."""
        pend = C.c_int(0)
        ns = (C.c_double * 3)()
        self._ck(self._L.gpt_trial_info(self._h, C.byref(pend), ns), "gpt_trial_info")
        return bool(pend.value), [ns[0], ns[1], ns[2]]

    TREE_NAMES = ("binary", "4-wide", "rebuilt", "split")

    def tree_info(self):
        """{in_use: name, pending, ns_per_ray: {name: ns}, inner_nodes: {name: n}} over the four candidate trees (gpt_tree_info)"""
        use, pend = C.c_int(0), C.c_int(0)
        ns, nn = (C.c_double * 4)(), (C.c_int * 4)()
        self._ck(self._L.gpt_tree_info(self._h, C.byref(use), C.byref(pend), ns, nn), "gpt_tree_info")
        return {"in_use": self.TREE_NAMES[use.value], "pending": bool(pend.value), "ns_per_ray": dict(zip(self.TREE_NAMES, list(ns))),
                "inner_nodes": dict(zip(self.TREE_NAMES, list(nn)))}

    def megakernel_tree_info(self):
        """{in_use: name, pending, ns_per_ray: {name: ns}}: the megakernel's own tree trial (gpt_megakernel_tree_info)"""
        use, pend = C.c_int(0), C.c_int(0)
        ns = (C.c_double * 4)()
        self._ck(self._L.gpt_megakernel_tree_info(self._h, C.byref(use), C.byref(pend), ns), "gpt_megakernel_tree_info")
        return {"in_use": self.TREE_NAMES[use.value], "pending": bool(pend.value), "ns_per_ray": dict(zip(self.TREE_NAMES, list(ns)))}

    def rebuild_info(self):
        """the rebuilt tree (gpt_rebuild_info): {state: 1 in use / 0 not / -1 on trial, inner_nodes, depth, ns_per_ray, build_ms}"""
        st, nn, dp = C.c_int(0), C.c_int(0), C.c_int(0)
        ns, ms = C.c_double(0.0), C.c_float(0.0)
        self._ck(self._L.gpt_rebuild_info(self._h, C.byref(st), C.byref(nn), C.byref(dp), C.byref(ns), C.byref(ms)), "gpt_rebuild_info")
        return {"state": st.value, "inner_nodes": nn.value, "depth": dp.value, "ns_per_ray": ns.value, "build_ms": ms.value}

    def kernel_times(self):
        """{class name: (ms, launches)} since the last reset_stats (needs FLAG_TIME_KERNELS)."""
        t = KernelTimes()
        self._ck(self._L.gpt_get_kernel_times(self._h, C.byref(t)), "gpt_get_kernel_times")
        names = ["generate", "extend", "shade", "collect", "accumulate", "resolve", "other"]
        return {n: (t.ms[i], int(t.launches[i])) for i, n in enumerate(names)}

    def measure_peaks(self):
        """micro-benchmarked peaks of this device (gpt_measure_peaks): dict of gather GB/s per memory level and warp instructions/s"""
        pk = Peaks()
        self._ck(self._L.gpt_measure_peaks(self._h, C.byref(pk)), "gpt_measure_peaks")
        return {k: getattr(pk, k) for k, _ in Peaks._fields_}

    def reset_stats(self):
        self._ck(self._L.gpt_reset_stats(self._h), "gpt_reset_stats")


class BasicScene:
    """Headless mirror of the reference's BasicScene (BasicScene.hpp:21-60): owns kernelInfo, the
    accumulation buffer and the draw buffer, and exposes launchKernel(info) + the per-frame contract
    of BasicScene::run (BasicScene.cpp:395-404).  The GLFW/ImGui/GL parts are out of scope."""

    def __init__(self, width, height, nodes=None, tris=None, indices=None, spheres=None, device=0, flags=0):
        import torch
        if not torch.cuda.is_available():
            raise GptError("BasicScene needs a CUDA device (no CPU fallback)")
        self.width, self.height = int(width), int(height)
        self.torch = torch
        self.dev = torch.device("cuda", device)
        self.ctx = Context(device, flags)
        self.info = default_kernel_info(width, height)
        self.spheres = spheres if spheres is not None else default_spheres()
        self.ctx.upload_scene(nodes, tris, indices, self.spheres)
        # cudaMalloc dev_drawRes / accumBuffer, BasicScene.cpp:137-151
        self.accumBuffer = torch.zeros(self.height * self.width * 3, dtype=torch.float32, device=self.dev)
        self.drawRes = torch.zeros(self.height * self.width, dtype=torch.int32, device=self.dev)
        self.info.accumBuffer = self.accumBuffer.data_ptr()
        self.info.dev_drawRes = self.drawRes.data_ptr()
        self.frameNumber = 0

    def launchKernel(self, info):
        self.ctx.launch(info, self.torch.cuda.current_stream(self.dev).cuda_stream)

    def frame(self, reset=False):
        """One iteration of the loop body BasicScene.cpp:395-404 (minus GL)."""
        self.torch.cuda.current_stream(self.dev).synchronize()
        self.info.hash = hash64(self.frameNumber)
        self.info.constantPdf = 1 if (reset or self.frameNumber == 0) else self.info.constantPdf + 1
        self.launchKernel(self.info)
        self.frameNumber += 1

    def run(self, frames):
        for _ in range(frames):
            self.frame()
        self.torch.cuda.current_stream(self.dev).synchronize()
        return self.accumBuffer.view(self.height, self.width, 3), self.drawRes.view(self.height, self.width)
