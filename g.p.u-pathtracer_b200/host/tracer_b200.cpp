// tracer_b200.cpp -- the adapter a maintainer of the reference adds to GpuPathTracer/ : compiled INSTEAD of tracer.cu, linked with
// -lgpt_b200.  It defines the one member function tracer.cu defines for the host side, BasicScene::launchKernel
// (BasicScene.hpp:32, tracer.cu:405-415), on top of the C-ABI of include/gpt.h.  Nothing else of the reference changes: it keeps
// allocating accumBuffer (vec3) and dev_drawRes, keeps uploading the CudaBVH arrays and the spheres (BasicScene.cpp:137-151,
// 213-216, 294-313) and keeps blitting dev_drawRes into its GL texture.
// tests/test_abi.py::test_adapter_compiles_against_reference compiles this file against the reference's own BasicScene.hpp.
#include "BasicScene.hpp"
#include "gpt.h"
#include <cstdio>
#include <cstdlib>

static gpt_ctx* g_ctx = nullptr;

// Call where the reference rewrites or re-uploads the BVH arrays behind kernelInfo.bvhData (the library traverses a re-packed
// copy cached by pointer identity; see "POINTER-IDENTITY RULE" in gpt.h).  The reference as shipped uploads them once
// (BasicScene.cpp:294-313), so it never needs this.
void gpt_b200_scene_changed() { if (g_ctx) gpt_invalidate_scene(g_ctx); }

void BasicScene::launchKernel(const kernelInfo& info)
{
    static_assert(sizeof(kernelInfo) == sizeof(gpt_kernel_info), "kernelInfo ABI drifted");
    if (!g_ctx && gpt_create(&g_ctx, 0) != GPT_OK) {
        fprintf(stderr, "gpt_create: %s\n", gpt_last_error_string(nullptr));
        exit(EXIT_FAILURE);                      // the reference's own error policy (utilfun.hpp:81-90)
    }
    // same call contract: asynchronous on stream 0, the caller syncs before touching `info` again (BasicScene.cpp:395)
    if (gpt_launch(g_ctx, reinterpret_cast<const gpt_kernel_info*>(&info), /*stream*/ nullptr) != GPT_OK) {
        fprintf(stderr, "gpt_launch: %s\n", gpt_last_error_string(g_ctx));
        exit(EXIT_FAILURE);
    }
}
