// gpt_headless.cpp -- headless replacement of the reference's viewer executable (App/main.cpp:7-20 ->
// BasicScene::BasicScene + BasicScene::run, BasicScene.cpp:65-318,368-477), C++ over the C-ABI of include/gpt.h.
//
// What it keeps of the reference's host side: the scene-setup block (default spheres, camera, kernelInfo defaults,
// OBJ -> indexed mesh -> BVH -> upload, BasicScene.cpp:154-318) and the per-frame contract (hash(frame), N = frames
// since reset, launch; :395-404).  What it drops: GLFW/ImGui/GL (no graphics stack on this box): the ImGui sliders
// become command-line flags that write the same kernelInfo fields (BasicScene.cpp:598-645), the GL blit becomes a PNG
// (+ optional PFM) file.  The BVH comes from the device-side builder (gpt_build_bvh); the reference's host
// SplitBVHBuilder output can be fed instead through gpt_upload_scene by an embedding application.
//
//   gpt_headless --obj scene.obj [--width 1280 --height 720 --spp 64 --depth 4 --mat diff|metal|spec|refr]
//                [--col r g b --emi r g b --bk r g b --phong 30 --air 1.0 --glass 1.4 --no-cull]
//                [--cam x y z yaw pitch] [--dist d] [--seed 0] [--mtl] [--checkpoint f] [--resume f] [--legacy] [--frames-per-pass 16] [--png out.png] [--pfm out.pfm] [--exr out.exr]
#include "../../include/gpt.h"
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

// the only CUDA-runtime entry points the host needs; resolved from the static runtime inside libgpt_b200.so is not
// possible, so device buffers are handled through the C-ABI helpers below (gpt_render_frame_host owns them)
static int die(gpt_ctx* ctx, const char* what, int rc)
{
    fprintf(stderr, "gpt_headless: %s failed (%d): %s\n", what, rc, gpt_last_error_string(ctx));
    if (ctx) gpt_destroy(ctx);
    return 1;
}

int main(int argc, char** argv)
{
    std::string obj, png = "render.png", pfm, exr, checkpoint, resume;
    bool use_mtl = false;
    int width = 1280, height = 720, device = 0;              // App/main.cpp:12
    unsigned spp = 16, depth = 4, per_pass = 16;
    unsigned long long seed = 0;
    int mat = GPT_MAT_DIFF, cull = 1;
    float col[3] = {246.0f / 256.0f, 246.0f / 255.0f, 70.0f / 255.0f}, emi[3] = {0, 0, 0}, bk[3] = {1, 1, 1};   // BasicScene.cpp:230-236
    float phong = 30.0f, air = 1.0f, glass = 1.4f;
    float cam_pos[3] = {0, 0, 0}, yaw = 0.0f, pitch = 0.0f, dist = -1.0f;
    bool have_cam = false;
    auto need = [&](int i, int n) { if (i + n >= argc) { fprintf(stderr, "gpt_headless: %s needs %d value(s)\n", argv[i], n); exit(2); } };
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        if (a == "--obj") { need(i, 1); obj = argv[++i]; }
        else if (a == "--png") { need(i, 1); png = argv[++i]; }
        else if (a == "--pfm") { need(i, 1); pfm = argv[++i]; }
        else if (a == "--exr") { need(i, 1); exr = argv[++i]; }
        else if (a == "--checkpoint") { need(i, 1); checkpoint = argv[++i]; }   // write float4 sums + sample count at the end
        else if (a == "--resume") { need(i, 1); resume = argv[++i]; }           // continue a checkpoint (same size and seed)
        else if (a == "--mtl") use_mtl = true;                                  // per-triangle materials from the OBJ's mtllib (N4)
        else if (a == "--width") { need(i, 1); width = atoi(argv[++i]); }
        else if (a == "--height") { need(i, 1); height = atoi(argv[++i]); }
        else if (a == "--device") { need(i, 1); device = atoi(argv[++i]); }
        else if (a == "--spp") { need(i, 1); spp = (unsigned)atoi(argv[++i]); }
        else if (a == "--depth") { need(i, 1); depth = (unsigned)atoi(argv[++i]); }
        else if (a == "--frames-per-pass") { need(i, 1); per_pass = (unsigned)atoi(argv[++i]); }
        else if (a == "--seed") { need(i, 1); seed = strtoull(argv[++i], nullptr, 10); }
        else if (a == "--phong") { need(i, 1); phong = (float)atof(argv[++i]); }
        else if (a == "--air") { need(i, 1); air = (float)atof(argv[++i]); }
        else if (a == "--glass") { need(i, 1); glass = (float)atof(argv[++i]); }
        else if (a == "--dist") { need(i, 1); dist = (float)atof(argv[++i]); }     // cam.dist; default height/60 (INT division, BasicScene.cpp:241)
        else if (a == "--no-cull") cull = 0;
        else if (a == "--mat") {
            need(i, 1); const std::string m = argv[++i];
            mat = m == "metal" ? GPT_MAT_METAL : m == "spec" ? GPT_MAT_SPEC : m == "refr" ? GPT_MAT_REFR : GPT_MAT_DIFF;
        }
        else if (a == "--col") { need(i, 3); for (int k = 0; k < 3; ++k) col[k] = (float)atof(argv[++i]); }
        else if (a == "--emi") { need(i, 3); for (int k = 0; k < 3; ++k) emi[k] = (float)atof(argv[++i]); }
        else if (a == "--bk") { need(i, 3); for (int k = 0; k < 3; ++k) bk[k] = (float)atof(argv[++i]); }
        else if (a == "--cam") { need(i, 5); for (int k = 0; k < 3; ++k) cam_pos[k] = (float)atof(argv[++i]); yaw = (float)atof(argv[++i]); pitch = (float)atof(argv[++i]); have_cam = true; }
        else { fprintf(stderr, "gpt_headless: unknown flag %s\n", a.c_str()); return 2; }
    }
    if (width < 2 || height < 2 || spp == 0) { fprintf(stderr, "gpt_headless: bad size / spp\n"); return 2; }
    if (per_pass == 0) per_pass = 1;

    gpt_ctx* ctx = nullptr;
    int rc = gpt_create(&ctx, device);
    if (rc) return die(nullptr, "gpt_create", rc);

    // ---- scene setup, BasicScene.cpp:177-318 ----
    gpt_sphere spheres[16]; size_t n_sph = 0;
    gpt_default_spheres(spheres, 16, &n_sph);
    gpt_kernel_info info;
    gpt_default_kernel_info(&info, width, height);
    info.depth = depth; info.triCurrentMat = mat; info.cullBackFaces = (uint8_t)cull;
    info.phongExpo = phong; info.air_ref_index = air; info.glass_ref_index = glass;
    for (int k = 0; k < 3; ++k) { info.col[k] = col[k]; info.emi[k] = emi[k]; info.bkColor[k] = bk[k]; }
    if (have_cam) gpt_set_camera(&info.cam, cam_pos, yaw, pitch);
    if (dist > 0.0f) info.cam.dist = dist;

    gpt_built_bvh bvh; memset(&bvh, 0, sizeof(bvh));
    if (!obj.empty()) {
        float* verts = nullptr; int32_t* tris = nullptr; size_t nv = 0, nt = 0;
        int32_t* face_mat = nullptr; gpt_tri_material* mats = nullptr; size_t n_mats = 0;
        const auto t0 = std::chrono::steady_clock::now();
        rc = use_mtl ? gpt_load_obj_materials(obj.c_str(), &verts, &nv, &tris, &nt, &face_mat, &mats, &n_mats)
                     : gpt_load_obj(obj.c_str(), &verts, &nv, &tris, &nt);
        if (rc) { fprintf(stderr, "gpt_headless: cannot read %s\n", obj.c_str()); gpt_destroy(ctx); return 1; }
        const double load_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        rc = gpt_build_bvh(ctx, verts, nv, tris, nt, 0, 8, &bvh);
        gpt_free_host(verts); gpt_free_host(tris);
        if (rc) return die(ctx, "gpt_build_bvh", rc);
        if (use_mtl && n_mats) {
            rc = gpt_set_triangle_materials(ctx, mats, n_mats, face_mat, nt);
            if (rc) return die(ctx, "gpt_set_triangle_materials", rc);
            printf("materials: %zu from the mtllib\n", n_mats);
        }
        gpt_free_host(face_mat); gpt_free_host(mats);
        printf("mesh: %zu vertices, %zu triangles (OBJ read %.2f s); BVH: %u inner nodes, %u leaves, depth %u, built on the device in %.2f ms\n",
               nv, nt, load_s, bvh.n_inner_nodes, bvh.n_leaves, bvh.depth, bvh.build_ms);
        rc = gpt_upload_scene(ctx, bvh.bvh.dev_triNode, bvh.bvh.triNodeSize, bvh.bvh.dev_triPtr, bvh.bvh.triSize,
                              bvh.bvh.dev_triIndicesPtr, bvh.bvh.triIndicesSize, spheres, n_sph, nullptr, 0, 1);
    } else {
        rc = gpt_upload_scene(ctx, nullptr, 0, nullptr, 0, nullptr, 0, spheres, n_sph, nullptr, 0, 0);
    }
    if (rc) return die(ctx, "gpt_upload_scene", rc);

    // ---- progressive loop, BasicScene.cpp:368-477 minus GL: each pass adds `per_pass` samples with global
    //      indices [done, done+n) into the library's float4 sum and returns the resolved frame to the host ----
    std::vector<uint32_t> rgba((size_t)width * height);
    std::vector<float> accum((size_t)width * height * 4), sum((size_t)width * height * 4, 0.0f);
    unsigned done = 0;
    if (!resume.empty()) {
        uint64_t n0 = 0, s0 = 0;
        rc = gpt_load_accum(resume.c_str(), sum.data(), width, height, &n0, &s0);
        if (rc || s0 != seed) { fprintf(stderr, "gpt_headless: cannot resume from %s (rc %d; size or seed mismatch?)\n", resume.c_str(), rc); gpt_destroy(ctx); return 1; }
        done = (unsigned)n0;
        spp += done;                      // --spp counts the samples to ADD; global indices continue at `done`
        printf("resumed %s at %u samples\n", resume.c_str(), done);
    }
    const auto t0 = std::chrono::steady_clock::now();
    while (done < spp) {
        const unsigned n = spp - done < per_pass ? spp - done : per_pass;
        rc = gpt_render_frame_host(ctx, &info, seed, done, n, rgba.data(), accum.data());
        if (rc) return die(ctx, "gpt_render_frame_host", rc);
        for (size_t i = 0; i < sum.size(); ++i) sum[i] += accum[i];
        done += n;
    }
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    gpt_stats st; gpt_get_stats(ctx, &st);
    printf("%ux%u, %u spp, depth %u: %.3f s, %.1f Mrays/s (%llu rays, %llu kernel launches)\n", width, height, spp, depth, secs,
           (double)st.rays / secs / 1e6, (unsigned long long)st.rays, (unsigned long long)st.kernel_launches);

    if (!checkpoint.empty()) {
        rc = gpt_save_accum(checkpoint.c_str(), sum.data(), width, height, spp, seed);
        if (rc) { fprintf(stderr, "gpt_headless: cannot write %s\n", checkpoint.c_str()); gpt_destroy(ctx); return 1; }
        printf("wrote %s (%u samples)\n", checkpoint.c_str(), spp);
    }
    // ---- resolve on the host from the summed float4 buffer: mean, clamp, pack (tracer.cu:386-398 semantics) ----
    std::vector<float> rgb((size_t)width * height * 3);
    for (size_t i = 0; i < (size_t)width * height; ++i) {
        const float inv = sum[4 * i + 3] > 0.0f ? 1.0f / sum[4 * i + 3] : 0.0f;
        uint32_t px = 0;
        for (int k = 0; k < 3; ++k) {
            const float m = sum[4 * i + k] * inv;
            rgb[3 * i + k] = m;
            const float c = m < 0.0f ? 0.0f : (m > 1.0f ? 1.0f : m);
            px |= (uint32_t)(unsigned char)(int)(255.0f * c) << (8 * k);
        }
        rgba[i] = px;
    }
    rc = gpt_write_png(png.c_str(), rgba.data(), width, height);
    if (rc) { fprintf(stderr, "gpt_headless: cannot write %s\n", png.c_str()); gpt_destroy(ctx); return 1; }
    printf("wrote %s\n", png.c_str());
    if (!pfm.empty()) {
        rc = gpt_write_pfm(pfm.c_str(), rgb.data(), width, height);
        if (rc) { fprintf(stderr, "gpt_headless: cannot write %s\n", pfm.c_str()); gpt_destroy(ctx); return 1; }
        printf("wrote %s\n", pfm.c_str());
    }
    if (!exr.empty()) {
        rc = gpt_write_exr(exr.c_str(), rgb.data(), width, height);
        if (rc) { fprintf(stderr, "gpt_headless: cannot write %s\n", exr.c_str()); gpt_destroy(ctx); return 1; }
        printf("wrote %s\n", exr.c_str());
    }
    if (bvh.bvh.dev_triNode) gpt_free_bvh(ctx, &bvh);
    gpt_destroy(ctx);
    return 0;
}
