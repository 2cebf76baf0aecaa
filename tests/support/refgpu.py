"""TEST-ONLY access to the REFERENCE's own CUDA kernel (never used by the product).

oracle/_ref/reftrace.cubin is `__global__ trace(const kernelInfo)` (GpuPathTracer/tracer.cu:343-400) compiled from
the reference sources where they lie by oracle/Makefile (PTX-level `vote.sync` fix).  It is loaded
here with the CUDA driver API into the primary context torch already uses, and launched exactly as
BasicScene::launchKernel does (tracer.cu:405-415): grid ((W+bx)/bx, (H+by)/by), block info.blockSize, the 296-byte
kernelInfo by value, default or given stream.  The product's ctypes `kernelInfo` is bit-compatible, so the same
struct instance drives both sides.

Uses: image parity of the CUDA path against the reference GPU tracer (tests/test_gpu_refkernel.py) and the
"reference kernel on this B200" baseline line of bench.py.
"""
import ctypes as C
import os

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
CUBIN = os.path.join(ROOT, "oracle", "_ref", "reftrace.cubin")
ENTRY = b"_Z5trace10kernelInfo"


def available():
    return os.path.exists(CUBIN)


class RefGpuKernel:
    def __init__(self, device=0):
        import torch
        self.torch = torch
        self.dev = torch.device("cuda", device)
        torch.zeros(1, device=self.dev)            # primary context created and current on this thread
        torch.cuda.set_device(self.dev)
        cu = C.CDLL("libcuda.so.1")
        self.cu = cu
        cu.cuModuleLoadData.argtypes = [C.POINTER(C.c_void_p), C.c_void_p]
        cu.cuModuleGetFunction.argtypes = [C.POINTER(C.c_void_p), C.c_void_p, C.c_char_p]
        cu.cuLaunchKernel.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_uint,
                                      C.c_uint, C.c_void_p, C.POINTER(C.c_void_p), C.c_void_p]
        cu.cuModuleUnload.argtypes = [C.c_void_p]
        self._ck(cu.cuInit(0), "cuInit")
        with open(CUBIN, "rb") as fh:
            self._image = fh.read()
        self.mod = C.c_void_p()
        self._ck(cu.cuModuleLoadData(C.byref(self.mod), self._image), "cuModuleLoadData")
        self.fn = C.c_void_p()
        self._ck(cu.cuModuleGetFunction(C.byref(self.fn), self.mod, ENTRY), "cuModuleGetFunction")

    def _ck(self, rc, what):
        if rc != 0:
            raise RuntimeError("%s failed: CUresult %d" % (what, rc))

    def launch(self, info, stream=None):
        """BasicScene::launchKernel(info), tracer.cu:405-415 (note the reference's non-ceil grid formula)."""
        bx, by = int(info.blockSize[0]), int(info.blockSize[1])
        gx, gy = (info.width + bx) // bx, (info.height + by) // by
        params = (C.c_void_p * 1)(C.addressof(info))
        self._ck(self.cu.cuLaunchKernel(self.fn, gx, gy, 1, bx, by, 1, 0, C.c_void_p(stream or 0), params, None), "cuLaunchKernel")

    def close(self):
        if getattr(self, "mod", None):
            self.torch.cuda.synchronize(self.dev)
            self.cu.cuModuleUnload(self.mod)
            self.mod = None


class RefScene:
    """The reference's device-side scene exactly as BasicScene.cpp:137-151,213-216,294-313 uploads it: the three
    CudaBVH arrays verbatim (nodes, debug triangles, indices), the Sphere array, a vec3 accumulation buffer and the
    RGBA8 draw buffer; fills the pointers of a kernelInfo."""

    def __init__(self, torch, info, scene, spheres, device=0):
        import numpy as np
        dev = torch.device("cuda", device)
        self.keep = []

        def up(a):
            t = torch.from_numpy(np.ascontiguousarray(a)).to(dev)
            self.keep.append(t)
            return t
        if scene is not None:
            n, t, i = up(scene["nodes"]), up(scene["tris"]), up(scene["indices"])
            info.bvhData.dev_triNode, info.bvhData.triNodeSize = n.data_ptr(), scene["nodes"].shape[0]
            info.bvhData.dev_triPtr, info.bvhData.triSize = t.data_ptr(), scene["tris"].shape[0]
            info.bvhData.dev_triIndicesPtr, info.bvhData.triIndicesSize = i.data_ptr(), scene["indices"].shape[0]
        ns = len(spheres)
        sp = up(np.frombuffer(bytes(spheres), dtype=np.uint8).copy())
        info.sphereTex, info.numSpheres = sp.data_ptr(), ns
        self.accum = torch.zeros(info.height * info.width * 3, dtype=torch.float32, device=dev)
        self.draw = torch.zeros(info.height * info.width, dtype=torch.int32, device=dev)
        info.accumBuffer, info.dev_drawRes = self.accum.data_ptr(), self.draw.data_ptr()
        info.blockSize[0], info.blockSize[1], info.blockSize[2] = 16, 16, 1      # BasicScene.cpp:224
        self.info = info
