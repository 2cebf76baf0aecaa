"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/gpt.h
declares, mirrors the reference's struct layouts, restates its scene-setup constants, and FAILS LOUDLY
without a GPU (no fallback)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from support import reflib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/GpuPathTracer"


def test_library_exports_every_declared_symbol(gpt):
    hdr = open(os.path.join(ROOT, "include", "gpt.h")).read()
    declared = sorted(set(re.findall(r"GPT_API[^;(]*?\b(gpt_[a-z0-9_]+)\s*\(", hdr)))
    assert len(declared) >= 19
    assert sorted(gpt.EXPORTS) == declared
    L = gpt.load_library()
    for name in declared:
        assert hasattr(L, name), name
    out = subprocess.check_output(["nm", "-D", "--defined-only", gpt.LIB_PATH]).decode()
    exported = set(re.findall(r" T (\w+)", out))
    assert set(declared) <= exported
    assert not any(("orc_" in s or "oracle" in s) for s in exported), "product must not contain oracle symbols"


def test_library_is_sm100a_cuda(gpt):
    out = subprocess.run(["cuobjdump", "-lelf", gpt.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out, out
    assert not re.search(r"sm_(?!100a)\d+", out), "only sm_100a cubins are shipped"


def test_struct_sizes_match_reference_probe(gpt):
    # SURVEY 8-a12 [probe]: Sphere 44, Box 52, CamInfo 84, BVHData 48, kernelInfo 296; offsets 32/48/184/268
    assert C.sizeof(gpt.Sphere) == 44 and C.sizeof(gpt.Box) == 52 and C.sizeof(gpt.CamInfo) == 84
    assert C.sizeof(gpt.BVHData) == 48 and C.sizeof(gpt.kernelInfo) == 296
    K = gpt.kernelInfo
    assert (K.accumBuffer.offset, K.cam.offset, K.bvhData.offset, K.emi.offset) == (32, 48, 184, 268)


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference not mounted")
def test_layout_against_reference_headers(tmp_path):
    """Compile offsetof/sizeof static_asserts against the REFERENCE's own CpuStructs.hpp / CommomStructs.hpp."""
    fields = ["blockSize", "hash", "dev_drawRes", "accumBuffer", "width", "height", "cam", "triangleTex", "numVerts", "sphereTex",
              "numSpheres", "cullBackFaces", "depth", "constantPdf", "bvhData", "time_elapsed", "air_ref_index", "glass_ref_index",
              "triCurrentMat", "clearBuffer", "bkColor", "phongExpo", "emi", "col"]
    cam = ["front", "right", "up", "pos", "dist", "pitch", "yaw", "aspect", "fov", "dirty", "bias", "enabled"]
    src = ['#include <cuda_runtime.h>', '#include <cstring>', '#include <cstdint>', '#include <glm/glm.hpp>',
           '#include "CommomStructs.hpp"', '#include "CpuStructs.hpp"', '#include "gpt.h"']
    src += ['static_assert(offsetof(kernelInfo,%s)==offsetof(gpt_kernel_info,%s),"%s");' % (f, f, f) for f in fields]
    src += ['static_assert(offsetof(CamInfo,%s)==offsetof(gpt_cam_info,%s),"%s");' % (f, f, f) for f in cam]
    src += ['static_assert(sizeof(kernelInfo)==sizeof(gpt_kernel_info),"");', 'static_assert(sizeof(CamInfo)==sizeof(gpt_cam_info),"");',
            'static_assert(sizeof(BVHData)==sizeof(gpt_bvh_data),"");', 'static_assert(sizeof(Sphere)==sizeof(gpt_sphere),"");',
            'static_assert(sizeof(Box)==sizeof(gpt_box),"");',
            'static_assert(offsetof(BVHData,dev_triIndicesPtr)==offsetof(gpt_bvh_data,dev_triIndicesPtr),"");',
            'static_assert(offsetof(BVHData,dev_triPtr)==offsetof(gpt_bvh_data,dev_triPtr),"");',
            'static_assert((int)Mat::DIFF==GPT_MAT_DIFF && (int)Mat::METAL==GPT_MAT_METAL && (int)Mat::SPEC==GPT_MAT_SPEC && (int)Mat::REFR==GPT_MAT_REFR,"");',
            'static_assert((int)GeoType::TRI==GPT_GEO_TRI && (int)GeoType::SPHERE==GPT_GEO_SPHERE && (int)GeoType::BOX==GPT_GEO_BOX && (int)GeoType::NONE==GPT_GEO_NONE,"");',
            'int main(){return 0;}']
    f = tmp_path / "abi.cpp"
    f.write_text("\n".join(src))
    subprocess.check_call(["/usr/bin/g++", "-std=c++11", "-w", "-fsyntax-only", "-D__host__=", "-D__device__=", "-I" + REF,
                           "-I/root/reference/deps/glm", "-I/usr/local/cuda/include", "-I" + os.path.join(ROOT, "include"), str(f)])


def test_hash64_matches_oracle(gpt):
    L = reflib.oracle()
    for k in [0, 1, 2, 3, 1000, 2 ** 33 + 5, 2 ** 64 - 1]:
        assert gpt.hash64(k) == L.orc_hash64(C.c_uint64(k))


def test_default_scene_constants(gpt):
    # BasicScene.cpp:177-259
    s = gpt.default_spheres()
    assert len(s) == 8
    assert list(s[0].posRad) == [0.0, -615.0, -20.0, 600.0] and s[0].mat == gpt.DIFF
    assert list(s[4].posRad) == [0.0, 0.0, -920.0, 600.0]
    assert list(s[6].posRad) == [13.0, -8.0, -35.0, 6.0] and s[6].mat == gpt.SPEC
    assert list(s[7].emi) == [0.0, 1.0, 1.0]
    np.testing.assert_allclose(list(s[2].col), [165 / 255, 15 / 255, 0.0], rtol=1e-7)
    for (w, h, dist) in [(1280, 720, 12.0), (1920, 1080, 18.0), (512, 512, 8.0), (3840, 2160, 36.0), (100, 100, 1.0)]:
        i = gpt.default_kernel_info(w, h)
        assert i.cam.dist == dist                           # height/60 INTEGER division
        assert i.cam.aspect == np.float32(w * 1.0 / h)
        assert abs(i.cam.fov - 1.0) < 1e-6
        assert i.depth == 4 and i.cullBackFaces == 1 and list(i.bkColor) == [1.0, 1.0, 1.0]
        assert list(i.blockSize) == [16, 16, 1] and i.triCurrentMat == gpt.DIFF and i.phongExpo == 30.0
        assert i.glass_ref_index == np.float32(1.4) and i.air_ref_index == 1.0
        np.testing.assert_allclose(list(i.col), [246 / 256, 246 / 255, 70 / 255], rtol=1e-7)
        np.testing.assert_allclose(list(i.cam.front), [0, 0, -1], atol=1e-7)
        np.testing.assert_allclose(list(i.cam.right), [1, 0, 0], atol=1e-7)
        np.testing.assert_allclose(list(i.cam.up), [0, 1, 0], atol=1e-7)


def test_set_camera_yaw_pitch(gpt):
    i = gpt.default_kernel_info(640, 360)
    gpt.set_camera(i, (1, 2, 3), yaw=90.0, pitch=30.0)
    f = np.array(list(i.cam.front))
    np.testing.assert_allclose(f, [np.cos(0) * np.cos(np.radians(30)), np.sin(np.radians(30)), np.sin(0) * np.cos(np.radians(30))], atol=1e-6)
    r = np.array(list(i.cam.right))
    np.testing.assert_allclose(r, [-f[2], 0, f[0]], atol=1e-7)          # not normalised when pitch != 0
    u = np.cross(r, f)
    np.testing.assert_allclose(list(i.cam.up), u / np.linalg.norm(u), atol=1e-6)
    assert list(i.cam.pos) == [1.0, 2.0, 3.0]
    gpt.set_camera(i, (0, 0, 0), yaw=0.0, pitch=120.0)
    assert i.cam.pitch == 89.0                                          # clamp, BasicScene.cpp:50-53


def test_no_gpu_fails_loudly(gpt):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(gpt.GptError) as e:
        gpt.Context(0)
    assert "no CUDA device" in str(e.value) or "CUDA" in str(e.value)
    with pytest.raises(gpt.GptError):
        gpt.BasicScene(64, 64)


def test_product_does_not_reference_oracle():
    """The product path must never import, link or call anything under oracle/."""
    pkg = os.path.join(ROOT, "g.p.u-pathtracer_b200")
    for dp, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")) or fn == "Makefile":
                txt = open(os.path.join(dp, fn), errors="ignore").read()
                assert "liboracle" not in txt and "oracle/" not in txt and "orc_" not in txt, fn
    ldd = subprocess.check_output(["ldd", os.path.join(pkg, "libgpt_b200.so")]).decode()
    assert "oracle" not in ldd and "cpuref" not in ldd and "refbvh" not in ldd


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference not mounted")
def test_adapter_compiles_against_reference(gpt, tmp_path):
    """the adapter a maintainer adds (g.p.u-pathtracer_b200/host/tracer_b200.cpp) is compiled against
    the REFERENCE's own BasicScene.hpp (BasicScene.hpp:32 declares launchKernel) and linked with libgpt_b200.so: it defines
    BasicScene::launchKernel(const kernelInfo&) and every gpt_* symbol it uses resolves."""
    deps = "/root/reference/deps"
    obj, so = str(tmp_path / "tracer_b200.o"), str(tmp_path / "libadapter.so")
    inc = ["-I" + REF, "-I" + deps + "/glm", "-I" + deps + "/glad/include", "-I" + deps + "/glfw/include", "-I" + deps + "/imgui",
           "-I" + deps + "/tinyobjloader", "-I/usr/local/cuda/include", "-I" + os.path.join(ROOT, "include")]
    src = os.path.join(os.path.dirname(gpt.LIB_PATH), "host", "tracer_b200.cpp")
    subprocess.check_call(["g++", "-std=c++11", "-fPIC", "-w", "-c", src, "-o", obj] + inc)
    syms = subprocess.check_output(["nm", "-C", obj]).decode()
    assert re.search(r" T BasicScene::launchKernel\(kernelInfo const&\)", syms), syms
    assert " U gpt_launch" in syms and " U gpt_invalidate_scene" in syms
    lib_dir = os.path.dirname(gpt.LIB_PATH)
    subprocess.check_call(["g++", "-shared", "-o", so, obj, "-L" + lib_dir, "-lgpt_b200", "-Wl,--no-undefined", "-Wl,-rpath," + lib_dir])
