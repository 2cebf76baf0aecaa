"""ctypes access to the TEST-ONLY libraries under oracle/ (never used by the product).

* oracle/liboracle.so      -- scalar CPU restatement of the hot path (oracle/oracle.cpp)
* oracle/_ref/librefbvh.so -- the reference's own SplitBVHBuilder/CudaBVH compiled in place
                              from /root/reference (oracle/Makefile); absent on machines
                              that never had the reference mounted and no prebuilt copy.
"""
import ctypes as C
import os
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle.so")
REFBVH_SO = os.path.join(ROOT, "oracle", "_ref", "librefbvh.so")


class OrcSphere(C.Structure):
    _fields_ = [("posRad", C.c_float * 4), ("emi", C.c_float * 3), ("col", C.c_float * 3), ("mat", C.c_int)]


class OrcBox(C.Structure):
    _fields_ = [("mn", C.c_float * 3), ("mx", C.c_float * 3), ("emi", C.c_float * 3), ("col", C.c_float * 3), ("mat", C.c_int)]


class OrcCam(C.Structure):
    _fields_ = [("front", C.c_float * 3), ("right", C.c_float * 3), ("up", C.c_float * 3), ("pos", C.c_float * 3),
                ("dist", C.c_float), ("aspect", C.c_float), ("fov", C.c_float)]


class OrcScene(C.Structure):
    _fields_ = [("nodes", C.c_void_p), ("n_node_vec4", C.c_size_t),
                ("tris", C.c_void_p), ("n_tri_vec4", C.c_size_t),
                ("indices", C.c_void_p), ("n_indices", C.c_size_t),
                ("spheres", C.c_void_p), ("n_spheres", C.c_size_t),
                ("boxes", C.c_void_p), ("n_boxes", C.c_size_t),
                ("tri_materials", C.c_void_p), ("n_tri_materials", C.c_size_t),
                ("face_material", C.c_void_p), ("n_faces", C.c_size_t)]


class OrcParams(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("cam", OrcCam),
                ("cull_back_faces", C.c_int), ("depth", C.c_uint32),
                ("air_ref_index", C.c_float), ("glass_ref_index", C.c_float),
                ("tri_mat", C.c_int), ("bk_color", C.c_float * 3), ("phong_expo", C.c_float),
                ("emi", C.c_float * 3), ("col", C.c_float * 3), ("seed", C.c_uint64), ("flags", C.c_uint32)]


class OrcCounters(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("rays_by_bounce", C.c_uint64 * 16), ("inner_nodes", C.c_uint64),
                ("tri_tests", C.c_uint64), ("leaf_terms", C.c_uint64), ("tri_hits", C.c_uint64),
                ("sphere_tests", C.c_uint64), ("max_stack", C.c_uint64)]

    def algorithmic_bytes(self):
        """64*N_inner + 48*N_tri + 16*N_leaf + 4*[tri hit] + 44*N_sphere (per-pixel accum bytes added by caller)."""
        return 64 * self.inner_nodes + 48 * self.tri_tests + 16 * self.leaf_terms + 4 * self.tri_hits + 44 * self.sphere_tests


ORC_FLAG_NO_JITTER = 1
ORC_FLAG_METAL_FIXED = 2
ORC_FLAG_REF_RNG = 4          # replay the reference's cuRAND XORWOW stream + CUDA sinf/cosf (oracle.h)
ORC_FLAG_REF_RNG_SWAP = 8     # A/B: other argument order at cudaUtils.h call site tracer.cu:170

_oracle = None


def oracle():
    global _oracle
    if _oracle is None:
        L = C.CDLL(ORACLE_SO)
        L.orc_hash64.restype = C.c_uint64
        L.orc_hash64.argtypes = [C.c_uint64]
        L.orc_u01.restype = C.c_float
        L.orc_u01.argtypes = [C.c_uint32]
        L.orc_powf.restype = C.c_float
        L.orc_powf.argtypes = [C.c_float, C.c_float]
        L.orc_sincos2pi.argtypes = [C.c_float, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        L.orc_sphere_intersect.restype = C.c_float
        L.orc_box_intersect.restype = C.c_float
        _oracle = L
    return _oracle


def fptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class RefBvhResult(C.Structure):
    _fields_ = [("nodes", C.c_void_p), ("n_node_vec4", C.c_uint),
                ("tris", C.c_void_p), ("n_tri_vec4", C.c_uint),
                ("woop", C.c_void_p), ("n_woop_vec4", C.c_uint),
                ("indices", C.c_void_p), ("n_indices", C.c_uint),
                ("n_leaves", C.c_uint)]


def have_refbvh():
    return os.path.exists(REFBVH_SO)


def ref_build_bvh(verts, tris):
    """Run the reference's SplitBVHBuilder + CudaBVH(BVHLayout_Compact) on an indexed mesh."""
    L = C.CDLL(REFBVH_SO)
    verts = np.ascontiguousarray(verts, dtype=np.float32)
    tris = np.ascontiguousarray(tris, dtype=np.int32)
    r = RefBvhResult()
    rc = L.refbvh_build(fptr(verts), C.c_int(len(verts)), fptr(tris), C.c_int(len(tris)), C.byref(r))
    if rc != 0:
        raise RuntimeError("refbvh_build failed rc=%d" % rc)

    def grab(ptr, n, dt):
        buf = (C.c_char * (n * np.dtype(dt).itemsize)).from_address(ptr)
        return np.frombuffer(buf, dtype=dt).copy()

    out = dict(nodes=grab(r.nodes, r.n_node_vec4 * 4, np.float32).reshape(-1, 4),
               tris=grab(r.tris, r.n_tri_vec4 * 4, np.float32).reshape(-1, 4),
               woop=grab(r.woop, r.n_woop_vec4 * 4, np.float32).reshape(-1, 4),
               indices=grab(r.indices, r.n_indices, np.int32),
               n_leaves=int(r.n_leaves))
    L.refbvh_free(C.byref(r))
    return out
