// gpt_host.cpp -- host-side helpers of the C-ABI that need no GPU: the reference's scene-setup
// constants restated so a headless caller can fill gpt_kernel_info exactly as BasicScene's
// constructor does (BasicScene.cpp:177-259), the per-frame seed hash (utilfun.cpp:380-389) and the
// yaw/pitch camera basis (BasicScene.cpp:42-64).
#include "../../include/gpt.h"
#include <cmath>
#include <cstring>

extern "C" {

uint64_t gpt_hash64(uint64_t key)
{
    // 64-bit Thomas Wang mix, utilfun.cpp:380-389
    key = (~key) + (key << 21);
    key ^= key >> 24;
    key = (key + (key << 3)) + (key << 8);
    key ^= key >> 14;
    key = (key + (key << 2)) + (key << 4);
    key ^= key >> 28;
    key += key << 31;
    return key;
}

static gpt_sphere mk_sphere(float x, float y, float z, float r, float er, float eg, float eb,
                            float cr, float cg, float cb, gpt_mat m)
{
    gpt_sphere s;
    s.posRad[0] = x; s.posRad[1] = y; s.posRad[2] = z; s.posRad[3] = r;
    s.emi[0] = er; s.emi[1] = eg; s.emi[2] = eb;
    s.col[0] = cr; s.col[1] = cg; s.col[2] = cb;
    s.mat = (int32_t)m;
    return s;
}

int gpt_default_spheres(gpt_sphere* out, size_t cap, size_t* n)
{
    // BasicScene.cpp:180-202: six r=600 "walls" (all emissive), one mirror ball, one emissive ball
    const float rad = 600.0f, pushX = 20.0f, pushY = 15.0f;
    const gpt_sphere s[8] = {
        mk_sphere(0.0f, -pushY - rad, -20.0f, rad, 197.0f / 255.0f, 153.0f / 255.0f, 92.0f / 255.0f, 0.5f, 0.5f, 0.5f, GPT_MAT_DIFF),
        mk_sphere(0.0f, pushY + rad, -20.0f, rad, 197.0f / 255.0f, 153.0f / 255.0f, 92.0f / 255.0f, 0.1f, 0.3f, 0.4f, GPT_MAT_DIFF),
        mk_sphere(pushX + rad, 0.0f, -20.0f, rad, 165.0f / 255.0f, 15.0f / 255.0f, 0.0f, 165.0f / 255.0f, 15.0f / 255.0f, 0.0f, GPT_MAT_DIFF),
        mk_sphere(-pushX - rad, 0.0f, -20.0f, rad, 30.0f / 255.0f, 76.0f / 255.0f, (float)(14.0 / 255.0), 30.0f / 255.0f, 76.0f / 255.0f, (float)(14.0 / 255.0), GPT_MAT_DIFF),
        mk_sphere(0.0f, 0.0f, (float)(-rad * 1.5 - 20), rad, (float)(197.0f / 255.0), 153.0f / 255.0f, 92.0f / 255.0f, 1.0f, 1.0f, 1.0f, GPT_MAT_DIFF),
        mk_sphere(0.0f, 0.0f, (float)(rad * 1.5 + 20), rad, 0.0f, 1.0f, 0.8f, 0.5f, 0.5f, 0.5f, GPT_MAT_DIFF),
        mk_sphere(13.0f, -8.0f, -35.0f, 6.0f, 0.0f, 0.0f, 0.0f, 1.0f, 1.0f, 1.0f, GPT_MAT_SPEC),
        mk_sphere(10.0f, -15.0f, -68.0f, 10.0f, 0.0f, 1.0f, 1.0f, 1.0f, 1.0f, 1.0f, GPT_MAT_DIFF),
    };
    if (n) *n = 8;
    if (!out) return GPT_OK;
    if (cap < 8) return GPT_ERR_INVALID;
    memcpy(out, s, sizeof(s));
    return GPT_OK;
}

void gpt_set_camera(gpt_cam_info* cam, const float pos[3], float yaw_deg, float pitch_deg)
{
    // setPitchAndRoll / constructor, BasicScene.cpp:42-64, 245-255 (glm::radians = deg * pi/180 in float)
    const float k = 0.01745329251994329576923690768489f;
    if (pitch_deg > 89.0f) pitch_deg = 89.0f;
    if (pitch_deg < -89.0f) pitch_deg = -89.0f;
    cam->yaw = yaw_deg; cam->pitch = pitch_deg;
    cam->front[0] = cosf((yaw_deg - 90.0f) * k) * cosf(pitch_deg * k);
    cam->front[1] = sinf(pitch_deg * k);
    cam->front[2] = sinf((yaw_deg - 90.0f) * k) * cosf(pitch_deg * k);
    cam->right[0] = -cam->front[2]; cam->right[1] = 0.0f; cam->right[2] = cam->front[0];
    // up = normalize(cross(right, front))
    const float ux = cam->right[1] * cam->front[2] - cam->front[1] * cam->right[2];
    const float uy = cam->right[2] * cam->front[0] - cam->front[2] * cam->right[0];
    const float uz = cam->right[0] * cam->front[1] - cam->front[0] * cam->right[1];
    const float inv = 1.0f / sqrtf(ux * ux + uy * uy + uz * uz);
    cam->up[0] = ux * inv; cam->up[1] = uy * inv; cam->up[2] = uz * inv;
    if (pos) { cam->pos[0] = pos[0]; cam->pos[1] = pos[1]; cam->pos[2] = pos[2]; }
    cam->dirty = 1;
}

void gpt_default_kernel_info(gpt_kernel_info* info, int width, int height)
{
    memset(info, 0, sizeof(*info));
    // CpuStructs.hpp:45-72 member initialisers
    info->depth = 1; info->constantPdf = 1; info->cullBackFaces = 1;
    info->air_ref_index = 1.0f; info->glass_ref_index = 1.4f;
    info->triCurrentMat = GPT_MAT_DIFF; info->phongExpo = 30.0f;
    // BasicScene.cpp:220-236
    info->width = width; info->height = height;
    info->blockSize[0] = 16; info->blockSize[1] = 16; info->blockSize[2] = 1;
    info->depth = 4;
    info->bkColor[0] = info->bkColor[1] = info->bkColor[2] = 1.0f;
    info->emi[0] = info->emi[1] = info->emi[2] = 0.0f;
    info->col[0] = 246.0f / 256.0f; info->col[1] = (float)(246.0 / 255.0f); info->col[2] = (float)(70.0 / 255.0f);
    // BasicScene.cpp:239-259
    info->cam.dist = (float)(height / 60);            // INTEGER division
    info->cam.aspect = width * 1.0f / height;
    info->cam.fov = tanf(45.0f * 0.01745329251994329576923690768489f);
    info->cam.bias[0] = info->cam.bias[1] = 0.01f; info->cam.enabled = 1;
    const float origin[3] = {0.0f, 0.0f, 0.0f};
    gpt_set_camera(&info->cam, origin, 0.0f, 0.0f);
}

} // extern "C"
