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

// ---------------------------------------------------------------- headless scene / image I/O
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace {

// strtof on the "C" locale subset OBJ files use; returns false at end of line
bool next_float(const char*& p, float* out)
{
    while (*p == ' ' || *p == '\t') ++p;
    if (*p == '\0' || *p == '\n' || *p == '\r' || *p == '#') return false;
    char* end = nullptr;
    *out = strtof(p, &end);
    if (end == p) return false;
    p = end;
    return true;
}

uint32_t crc_table[256];
bool crc_ready = false;
uint32_t crc32_update(uint32_t c, const unsigned char* buf, size_t n)
{
    if (!crc_ready) {
        for (uint32_t i = 0; i < 256; ++i) { uint32_t k = i; for (int j = 0; j < 8; ++j) k = (k & 1u) ? 0xEDB88320u ^ (k >> 1) : k >> 1; crc_table[i] = k; }
        crc_ready = true;
    }
    for (size_t i = 0; i < n; ++i) c = crc_table[(c ^ buf[i]) & 0xffu] ^ (c >> 8);
    return c;
}

void put_be32(std::vector<unsigned char>& v, uint32_t x) { v.push_back(x >> 24); v.push_back(x >> 16); v.push_back(x >> 8); v.push_back(x); }

bool write_chunk(FILE* f, const char* tag, const std::vector<unsigned char>& data)
{
    std::vector<unsigned char> hdr; put_be32(hdr, (uint32_t)data.size());
    unsigned char t[4] = {(unsigned char)tag[0], (unsigned char)tag[1], (unsigned char)tag[2], (unsigned char)tag[3]};
    uint32_t c = crc32_update(0xffffffffu, t, 4);
    if (!data.empty()) c = crc32_update(c, data.data(), data.size());
    c ^= 0xffffffffu;
    std::vector<unsigned char> tail; put_be32(tail, c);
    return fwrite(hdr.data(), 1, 4, f) == 4 && fwrite(t, 1, 4, f) == 4 &&
           (data.empty() || fwrite(data.data(), 1, data.size(), f) == data.size()) && fwrite(tail.data(), 1, 4, f) == 4;
}

} // namespace

extern "C" {

// uf::loadIndexedTris (utilfun.cpp:454-524): every `v` in file order, per face the 0-based vertex index of each
// corner, polygons fan-triangulated (v0, v[k], v[k+1]) like tinyobjloader; normals / uvs / materials dropped
// (SURVEY F5).  Unlike the reference (assert(shapes.size()==1), :474) several o/g shapes are merged into one mesh.
static int load_obj_impl(const char* path, float** verts_out, size_t* n_verts, int32_t** tris_out, size_t* n_tris,
                         std::vector<int32_t>* face_mat, std::vector<std::string>* mat_names, std::string* mtllib)
{
    if (!path || !verts_out || !n_verts || !tris_out || !n_tris) return GPT_ERR_INVALID;
    *verts_out = nullptr; *tris_out = nullptr; *n_verts = 0; *n_tris = 0;
    FILE* f = fopen(path, "rb");
    if (!f) return GPT_ERR_IO;
    std::vector<float> v;
    std::vector<int32_t> t;
    int32_t cur_mat = -1;
    std::vector<long long> corner;
    std::string line;
    char buf[4096];
    bool bad = false;
    while (fgets(buf, sizeof(buf), f)) {
        line.assign(buf);
        while (!line.empty() && line.back() != '\n' && fgets(buf, sizeof(buf), f)) line.append(buf);   // long lines
        const char* p = line.c_str();
        while (*p == ' ' || *p == '\t') ++p;
        if (p[0] == 'v' && (p[1] == ' ' || p[1] == '\t')) {
            p += 2;
            float x[3] = {0.0f, 0.0f, 0.0f};
            for (int k = 0; k < 3; ++k) if (!next_float(p, &x[k])) { bad = true; break; }
            v.push_back(x[0]); v.push_back(x[1]); v.push_back(x[2]);
        } else if (p[0] == 'f' && (p[1] == ' ' || p[1] == '\t')) {
            p += 2;
            corner.clear();
            const long long nv = (long long)(v.size() / 3);
            while (true) {
                while (*p == ' ' || *p == '\t') ++p;
                if (*p == '\0' || *p == '\n' || *p == '\r' || *p == '#') break;
                char* end = nullptr;
                const long long i = strtoll(p, &end, 10);
                if (end == p) { bad = true; break; }
                p = end;
                while (*p != '\0' && *p != ' ' && *p != '\t' && *p != '\n' && *p != '\r') ++p;   // skip /vt/vn
                const long long idx = i > 0 ? i - 1 : nv + i;
                if (idx < 0 || idx >= nv) { bad = true; break; }
                corner.push_back(idx);
            }
            for (size_t k = 1; k + 1 < corner.size(); ++k) {
                t.push_back((int32_t)corner[0]); t.push_back((int32_t)corner[k]); t.push_back((int32_t)corner[k + 1]);
                if (face_mat) face_mat->push_back(cur_mat);
            }
        } else if (mat_names && (strncmp(p, "usemtl", 6) == 0 || strncmp(p, "mtllib", 6) == 0) && (p[6] == ' ' || p[6] == '\t')) {
            const bool lib = p[0] == 'm';
            p += 7;
            while (*p == ' ' || *p == '\t') ++p;
            std::string name(p);
            while (!name.empty() && (name.back() == '\n' || name.back() == '\r' || name.back() == ' ' || name.back() == '\t')) name.pop_back();
            if (lib) { if (mtllib) *mtllib = name; }
            else {
                cur_mat = -1;
                for (size_t k = 0; k < mat_names->size(); ++k) if ((*mat_names)[k] == name) cur_mat = (int32_t)k;
                if (cur_mat < 0) { mat_names->push_back(name); cur_mat = (int32_t)mat_names->size() - 1; }
            }
        }
        if (bad) break;
    }
    fclose(f);
    if (bad) return GPT_ERR_IO;
    float* vo = (float*)malloc(v.size() * sizeof(float) + 4);
    int32_t* to = (int32_t*)malloc(t.size() * sizeof(int32_t) + 4);
    if (!vo || !to) { free(vo); free(to); return GPT_ERR_INVALID; }
    if (!v.empty()) memcpy(vo, v.data(), v.size() * sizeof(float));
    if (!t.empty()) memcpy(to, t.data(), t.size() * sizeof(int32_t));
    *verts_out = vo; *n_verts = v.size() / 3; *tris_out = to; *n_tris = t.size() / 3;
    return GPT_OK;
}

int gpt_load_obj(const char* path, float** verts_out, size_t* n_verts, int32_t** tris_out, size_t* n_tris)
{
    return load_obj_impl(path, verts_out, n_verts, tris_out, n_tris, nullptr, nullptr, nullptr);
}

// OBJ + its .mtl (what tinyobj parses and the reference then drops, utilfun.cpp:457-497): per
// triangle the index of its `usemtl` material (-1 = none), and one gpt_tri_material per material name:
// col = Kd, emi = Ke, Mat from illum (0-2 DIFF, 3 METAL, 5 SPEC mirror, 4/6/7/9 REFR glass).  Materials used in the
// OBJ but missing from the .mtl (or no mtllib at all) get the reference's default triangle colour
// (BasicScene.cpp:234-236) and DIFF.
int gpt_load_obj_materials(const char* path, float** verts_out, size_t* n_verts, int32_t** tris_out, size_t* n_tris,
                           int32_t** face_material_out, gpt_tri_material** materials_out, size_t* n_materials)
{
    if (!face_material_out || !materials_out || !n_materials) return GPT_ERR_INVALID;
    *face_material_out = nullptr; *materials_out = nullptr; *n_materials = 0;
    std::vector<int32_t> face_mat; std::vector<std::string> names; std::string mtllib;
    int rc = load_obj_impl(path, verts_out, n_verts, tris_out, n_tris, &face_mat, &names, &mtllib);
    if (rc != GPT_OK) return rc;
    std::vector<gpt_tri_material> mats(names.size());
    for (gpt_tri_material& m : mats) { m.col[0] = 246.0f / 256.0f; m.col[1] = 246.0f / 255.0f; m.col[2] = 70.0f / 255.0f; m.emi[0] = m.emi[1] = m.emi[2] = 0.0f; m.mat = GPT_MAT_DIFF; }
    if (!mtllib.empty()) {
        std::string dir(path);
        const size_t slash = dir.find_last_of('/');
        dir = slash == std::string::npos ? std::string() : dir.substr(0, slash + 1);
        FILE* f = fopen((dir + mtllib).c_str(), "rb");
        if (f) {
            char buf[1024];
            int cur = -1;
            while (fgets(buf, sizeof(buf), f)) {
                const char* p = buf;
                while (*p == ' ' || *p == '\t') ++p;
                if (strncmp(p, "newmtl", 6) == 0 && (p[6] == ' ' || p[6] == '\t')) {
                    p += 7; while (*p == ' ' || *p == '\t') ++p;
                    std::string name(p);
                    while (!name.empty() && (name.back() == '\n' || name.back() == '\r' || name.back() == ' ' || name.back() == '\t')) name.pop_back();
                    cur = -1;
                    for (size_t k = 0; k < names.size(); ++k) if (names[k] == name) cur = (int)k;
                } else if (cur >= 0 && p[0] == 'K' && (p[1] == 'd' || p[1] == 'e') && (p[2] == ' ' || p[2] == '\t')) {
                    float* dst = p[1] == 'd' ? mats[cur].col : mats[cur].emi;
                    p += 3;
                    float x[3];
                    if (next_float(p, &x[0]) && next_float(p, &x[1]) && next_float(p, &x[2])) { dst[0] = x[0]; dst[1] = x[1]; dst[2] = x[2]; }
                } else if (cur >= 0 && strncmp(p, "illum", 5) == 0) {
                    const int il = atoi(p + 5);
                    mats[cur].mat = il == 3 ? GPT_MAT_METAL : il == 5 ? GPT_MAT_SPEC : (il == 4 || il == 6 || il == 7 || il == 9) ? GPT_MAT_REFR : GPT_MAT_DIFF;
                }
            }
            fclose(f);
        }
    }
    int32_t* fm = (int32_t*)malloc(face_mat.size() * sizeof(int32_t) + 4);
    gpt_tri_material* mo = (gpt_tri_material*)malloc(mats.size() * sizeof(gpt_tri_material) + 4);
    if (!fm || !mo) { free(fm); free(mo); free(*verts_out); free(*tris_out); *verts_out = nullptr; *tris_out = nullptr; return GPT_ERR_INVALID; }
    if (!face_mat.empty()) memcpy(fm, face_mat.data(), face_mat.size() * sizeof(int32_t));
    if (!mats.empty()) memcpy(mo, mats.data(), mats.size() * sizeof(gpt_tri_material));
    *face_material_out = fm; *materials_out = mo; *n_materials = mats.size();
    return GPT_OK;
}

void gpt_free_host(void* p) { free(p); }

// dev_drawRes layout (tracer.cu:394-398): W*H uint32 0x00BBGGRR, row 0 = BOTTOM of the image (getCamRayDir,
// cudaUtils.h:111-134) -> 8-bit RGB PNG, top row first.  Replaces the GL texture blit + quad draw
// (BasicScene.cpp:424-432,468) and lodepng in the CPU tracer (renderer.cpp:224-245).  Stored (uncompressed)
// deflate blocks: no zlib dependency.
int gpt_write_png(const char* path, const uint32_t* rgba8, int width, int height)
{
    if (!path || !rgba8 || width <= 0 || height <= 0) return GPT_ERR_INVALID;
    FILE* f = fopen(path, "wb");
    if (!f) return GPT_ERR_IO;
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    bool ok = fwrite(sig, 1, 8, f) == 8;
    std::vector<unsigned char> ihdr;
    put_be32(ihdr, (uint32_t)width); put_be32(ihdr, (uint32_t)height);
    ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);   // 8-bit RGB
    ok = ok && write_chunk(f, "IHDR", ihdr);
    const size_t stride = 1 + 3 * (size_t)width;
    std::vector<unsigned char> raw(stride * (size_t)height);
    for (int y = 0; y < height; ++y) {
        const uint32_t* src = rgba8 + (size_t)(height - 1 - y) * width;      // vertical flip
        unsigned char* dst = raw.data() + stride * (size_t)y;
        dst[0] = 0;                                                           // filter: none
        for (int x = 0; x < width; ++x) { const uint32_t c = src[x]; dst[1 + 3 * x] = c & 0xff; dst[2 + 3 * x] = (c >> 8) & 0xff; dst[3 + 3 * x] = (c >> 16) & 0xff; }
    }
    std::vector<unsigned char> z;
    z.reserve(raw.size() + raw.size() / 65535 * 5 + 16);
    z.push_back(0x78); z.push_back(0x01);
    uint32_t a = 1, b = 0;
    size_t pos = 0;
    while (pos < raw.size() || raw.empty()) {
        const size_t n = raw.size() - pos < 65535 ? raw.size() - pos : 65535;
        z.push_back(pos + n == raw.size() ? 1 : 0);
        z.push_back(n & 0xff); z.push_back(n >> 8); z.push_back(~n & 0xff); z.push_back((~n >> 8) & 0xff);
        z.insert(z.end(), raw.begin() + pos, raw.begin() + pos + n);
        pos += n;
        if (pos >= raw.size()) break;
    }
    for (size_t i = 0; i < raw.size();) {                                     // adler32, modulo deferred per 5552 bytes
        const size_t n = raw.size() - i < 5552 ? raw.size() - i : 5552;
        for (size_t k = 0; k < n; ++k) { a += raw[i + k]; b += a; }
        a %= 65521u; b %= 65521u; i += n;
    }
    put_be32(z, (b << 16) | a);
    ok = ok && write_chunk(f, "IDAT", z) && write_chunk(f, "IEND", std::vector<unsigned char>());
    ok = (fclose(f) == 0) && ok;
    return ok ? GPT_OK : GPT_ERR_IO;
}

// Linear float image (PFM "PF", little endian, bottom row first -- the accumulation buffer's own row order):
// rgb = W*H*3 floats as gpt_resolve's out_rgb_f32 writes them.
int gpt_write_pfm(const char* path, const float* rgb, int width, int height)
{
    if (!path || !rgb || width <= 0 || height <= 0) return GPT_ERR_INVALID;
    FILE* f = fopen(path, "wb");
    if (!f) return GPT_ERR_IO;
    bool ok = fprintf(f, "PF\n%d %d\n-1.0\n", width, height) > 0;
    const size_t n = 3 * (size_t)width * (size_t)height;
    ok = ok && fwrite(rgb, sizeof(float), n, f) == n;
    ok = (fclose(f) == 0) && ok;
    return ok ? GPT_OK : GPT_ERR_IO;
}

// Linear float image as OpenEXR 2.0 (north_star item 3: "headless PNG/EXR output"): single-part scan-line file, no compression,
// three 32-bit FLOAT channels (stored alphabetically: B, G, R), increasing-Y line order with the TOP row first -- so the bottom-up
// rows of out_rgb_f32 (row 0 = bottom, tracer.cu:352-358) are written last to first, the same flip gpt_write_png makes.
// Layout per the OpenEXR file-layout document: magic 20000630, version 2, header attributes (name, type, size, value), a zero
// byte, the table of 64-bit scan-line offsets, then per scan line: int32 y, int32 byte count, the channels' rows one after another.
namespace {
struct ExrBuf {
    std::vector<unsigned char> b;
    void raw(const void* p, size_t n) { const unsigned char* c = (const unsigned char*)p; b.insert(b.end(), c, c + n); }
    void str(const char* z) { raw(z, strlen(z) + 1); }
    void i32(int32_t v) { raw(&v, 4); }
    void f32(float v) { raw(&v, 4); }
    void attr(const char* name, const char* type, int32_t size) { str(name); str(type); i32(size); }
};
}

int gpt_write_exr(const char* path, const float* rgb, int width, int height)
{
    if (!path || !rgb || width <= 0 || height <= 0) return GPT_ERR_INVALID;
    ExrBuf h;
    h.i32(20000630); h.i32(2);
    h.attr("channels", "chlist", 3 * 18 + 1);
    for (const char* ch : {"B", "G", "R"}) { h.str(ch); h.i32(2 /* FLOAT */); h.i32(0 /* pLinear + reserved */); h.i32(1); h.i32(1); }
    h.b.push_back(0);
    h.attr("compression", "compression", 1); h.b.push_back(0 /* NO_COMPRESSION */);
    h.attr("dataWindow", "box2i", 16); h.i32(0); h.i32(0); h.i32(width - 1); h.i32(height - 1);
    h.attr("displayWindow", "box2i", 16); h.i32(0); h.i32(0); h.i32(width - 1); h.i32(height - 1);
    h.attr("lineOrder", "lineOrder", 1); h.b.push_back(0 /* INCREASING_Y */);
    h.attr("pixelAspectRatio", "float", 4); h.f32(1.0f);
    h.attr("screenWindowCenter", "v2f", 8); h.f32(0.0f); h.f32(0.0f);
    h.attr("screenWindowWidth", "float", 4); h.f32(1.0f);
    h.b.push_back(0);                                            // end of header
    const uint64_t line_bytes = 8 + 12 * (uint64_t)width;        // y, size, three float rows
    uint64_t off = h.b.size() + 8 * (uint64_t)height;
    for (int y = 0; y < height; ++y, off += line_bytes) h.raw(&off, 8);
    FILE* f = fopen(path, "wb");
    if (!f) return GPT_ERR_IO;
    bool ok = fwrite(h.b.data(), 1, h.b.size(), f) == h.b.size();
    std::vector<float> row(3 * (size_t)width);
    for (int y = 0; y < height && ok; ++y) {
        const float* src = rgb + 3 * (size_t)width * (size_t)(height - 1 - y);     // EXR y = 0 is the top row
        for (int x = 0; x < width; ++x) { row[x] = src[3 * x + 2]; row[width + x] = src[3 * x + 1]; row[2 * (size_t)width + x] = src[3 * x]; }
        const int32_t hdr[2] = {y, (int32_t)(12 * width)};
        ok = fwrite(hdr, 4, 2, f) == 2 && fwrite(row.data(), sizeof(float), row.size(), f) == row.size();
    }
    ok = (fclose(f) == 0) && ok;
    return ok ? GPT_OK : GPT_ERR_IO;
}

// Accumulation checkpoint / resume: the reference keeps accumBuffer + constantPdf in process memory
// only (BasicScene.cpp:399).  File = "GPTACC1\0", int32 width, height, uint64 samples done, uint64 seed, then W*H
// float4 (sum r, g, b, sample count) as gpt_render accumulates them.
int gpt_save_accum(const char* path, const float* accum_f4, int width, int height, uint64_t samples_done, uint64_t seed)
{
    if (!path || !accum_f4 || width <= 0 || height <= 0) return GPT_ERR_INVALID;
    FILE* f = fopen(path, "wb");
    if (!f) return GPT_ERR_IO;
    const int32_t wh[2] = {width, height};
    const uint64_t meta[2] = {samples_done, seed};
    const size_t n = 4 * (size_t)width * (size_t)height;
    bool ok = fwrite("GPTACC1", 1, 8, f) == 8 && fwrite(wh, 4, 2, f) == 2 && fwrite(meta, 8, 2, f) == 2 && fwrite(accum_f4, sizeof(float), n, f) == n;
    ok = (fclose(f) == 0) && ok;
    return ok ? GPT_OK : GPT_ERR_IO;
}

int gpt_load_accum(const char* path, float* accum_f4, int width, int height, uint64_t* samples_done, uint64_t* seed)
{
    if (!path || !accum_f4 || width <= 0 || height <= 0) return GPT_ERR_INVALID;
    FILE* f = fopen(path, "rb");
    if (!f) return GPT_ERR_IO;
    char magic[8]; int32_t wh[2]; uint64_t meta[2];
    const size_t n = 4 * (size_t)width * (size_t)height;
    bool ok = fread(magic, 1, 8, f) == 8 && memcmp(magic, "GPTACC1", 8) == 0 && fread(wh, 4, 2, f) == 2 && fread(meta, 8, 2, f) == 2;
    if (ok && (wh[0] != width || wh[1] != height)) { fclose(f); return GPT_ERR_INVALID; }
    ok = ok && fread(accum_f4, sizeof(float), n, f) == n;
    fclose(f);
    if (!ok) return GPT_ERR_IO;
    if (samples_done) *samples_done = meta[0];
    if (seed) *seed = meta[1];
    return GPT_OK;
}

} // extern "C"
