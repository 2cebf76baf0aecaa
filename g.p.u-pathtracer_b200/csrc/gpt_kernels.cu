// gpt_kernels.cu -- the render megakernel, test-hook kernels and resolve, with their launchers.
//
// k_render_mega replaces __global__ trace + getSample (tracer.cu:343-400, 27-339): instead of
// "1 thread = 1 pixel = 1 path per launch" on a W/16 x H/16 grid, a persistent grid (a multiple of the
// SM count) pulls pixel jobs from a global counter, one warp-aggregated atomic per refill.  A lane keeps
// a pixel for all samples of the launch, runs one bounce per loop trip, and refills the moment its pixel
// is finished, so lanes of a warp are always tracing (the reference idles every lane until the longest
// path of the warp ends).  Sums are kept in registers and written once per pixel: no atomics on the
// accumulation buffer and a deterministic summation order.
#include "gpt_shade.cuh"
#include "gpt_kernels.h"
#include <cstdlib>

#define MEGA_THREADS 128
#define TILE_W 8
#define TILE_H 4

struct SmemScene {
    float spheres[GPT_MAX_SMEM_SPHERES * 11];
    float boxes[GPT_MAX_SMEM_BOXES * 13];
    unsigned long long bounce_cnt[16];
};

__device__ __forceinline__ void stage_scene(const DevScene& sc, SmemScene& sm, const float*& s_spheres, const float*& s_boxes)
{
    s_spheres = sc.spheres; s_boxes = sc.boxes;
    if (sc.n_spheres <= GPT_MAX_SMEM_SPHERES) {
        for (int i = threadIdx.x; i < sc.n_spheres * 11; i += blockDim.x) sm.spheres[i] = sc.spheres[i];
        s_spheres = sm.spheres;
    }
    if (sc.n_boxes <= GPT_MAX_SMEM_BOXES) {
        for (int i = threadIdx.x; i < sc.n_boxes * 13; i += blockDim.x) sm.boxes[i] = sc.boxes[i];
        s_boxes = sm.boxes;
    }
    if (threadIdx.x < 16) sm.bounce_cnt[threadIdx.x] = 0ull;
    __syncthreads();
}

// tracer.cu:85-105 (clamp + rgbToUint) with the uchar truncation of :369,394-396
__device__ __forceinline__ uint32_t pack_rgb(float ax, float ay, float az)
{
    const unsigned char cx = (unsigned char)(int)(255.0f * ax), cy = (unsigned char)(int)(255.0f * ay), cz = (unsigned char)(int)(255.0f * az);
    const float r = fmaxf(0.0f, fminf((float)cx, 255.0f)), g = fmaxf(0.0f, fminf((float)cy, 255.0f)), b = fmaxf(0.0f, fminf((float)cz, 255.0f));
    return ((uint32_t)b << 16) | ((uint32_t)g << 8) | (uint32_t)r;
}
__device__ __forceinline__ float clamp01(float v) { return fminf(fmaxf(v, 0.0f), 1.0f); }

// LEGACY = false: accum_f4[pix] += (sum rgb, count)            (native contract)
// LEGACY = true : vec3 running mean + clamp + RGBA8 pack        (tracer.cu:386-398), sample_count must be 1
template <bool LEGACY>
__global__ void __launch_bounds__(MEGA_THREADS, 4)
k_render_mega(const DevScene sc, const RenderParams p, float4* __restrict__ accum_f4,
              float* __restrict__ accum_v3, uint32_t* __restrict__ draw, const unsigned long long constant_pdf,
              const unsigned long long sample_begin, const uint32_t sample_count,
              unsigned int* __restrict__ work_counter, unsigned long long* __restrict__ stats)
{
    __shared__ SmemScene sm;
    const float* s_spheres; const float* s_boxes;
    stage_scene(sc, sm, s_spheres, s_boxes);

    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int tiles_x = (p.width + TILE_W - 1) / TILE_W, tiles_y = (p.height + TILE_H - 1) / TILE_H;
    const unsigned total_jobs = (unsigned)tiles_x * (unsigned)tiles_y * 32u;
    const bool cull = p.cull != 0;
    const bool count_bounces = (p.flags & RP_COUNT_BOUNCES) != 0;

    bool have_job = false, done = false;
    int px = 0, py = 0; uint32_t pixel = 0, s = 0, bounce = 0;
    PathState ps;
    TravState ts;
    int stack[GPT_STACK_SIZE];
    f3 sum = mk3(0.0f, 0.0f, 0.0f);
    unsigned long long n_rays = 0, n_paths = 0;
    const bool has_mesh = sc.n_nodes > 0;
    ts.node = GPT_SENTINEL; ts.leaf = 0; ts.sp = 0; ts.t = GPT_F32_MAX; ts.tri = -1; ts.tri_cur = 0;

    // One trip = for every lane: finish the previous bounce (sphere sweep + shading), move on to the next
    // sample / pixel if the path ended, arm the next ray, then traverse it to completion (while-while).
    // Schedules that re-arm lanes mid-traversal were measured and lost (profiles/r01_schedules.md): in a
    // megakernel re-arming costs a full shading step.  The wavefront pipeline (gpt_wavefront.cu), where
    // re-arming is a 32-byte load, is the default for gpt_render; this kernel is the single-launch alternative.
    while (true) {
        // ---- finish the bounce whose traversal just ended: tracer.cu:98-296 ----
        if (have_job && bounce < p.depth) {
            const SceneHit h = finish_scene_hit(sc, s_spheres, s_boxes, ts.t, ts.tri, ps.o, ps.d);
            const bool cont = shade_bounce(sc, p, s_spheres, s_boxes, h, pixel, sample_begin + s, bounce, ps);
            ++bounce;
            if (!cont) bounce = p.depth;                     // miss: path over (ps.accu = bkColor)
        }
        if (have_job && bounce >= p.depth) {
            // path finished: result = ps.accu (tracer.cu:305, or bkColor on a miss :141)
            sum.x += ps.accu.x; sum.y += ps.accu.y; sum.z += ps.accu.z;
            ++s;
            if (s < sample_count) {
                bounce = 0;
                start_path(p, px, py, pixel, sample_begin + s, ps); ++n_paths;
            } else {
                if (LEGACY) {
                    // tracer.cu:386-398
                    const float nm1 = (float)(constant_pdf - 1ull), inv = 1.0f / (float)constant_pdf;
                    float* a = accum_v3 + 3 * (size_t)pixel;
                    const float ax = clamp01((a[0] * nm1 + sum.x) * inv);
                    const float ay = clamp01((a[1] * nm1 + sum.y) * inv);
                    const float az = clamp01((a[2] * nm1 + sum.z) * inv);
                    a[0] = ax; a[1] = ay; a[2] = az;
                    if (draw) draw[pixel] = pack_rgb(ax, ay, az);
                } else {
                    float4 a = accum_f4[pixel];
                    a.x += sum.x; a.y += sum.y; a.z += sum.z; a.w += (float)sample_count;
                    accum_f4[pixel] = a;
                }
                have_job = false;
            }
        }
        // ---- warp-level work fetch: one atomic for all lanes that need a pixel ----
        const unsigned need = __ballot_sync(GPT_FULL_MASK, !have_job && !done);
        if (need) {
            const int leader = __ffs(need) - 1;
            unsigned base = 0;
            if ((int)lane == leader) base = atomicAdd(work_counter, (unsigned)__popc(need));
            base = __shfl_sync(GPT_FULL_MASK, base, leader);
            if (!have_job && !done) {
                const unsigned j = base + (unsigned)__popc(need & lt_mask);
                if (j >= total_jobs) done = true;
                else {
                    const unsigned tile = j >> 5, in = j & 31u;
                    px = (int)(tile % (unsigned)tiles_x) * TILE_W + (int)(in & (TILE_W - 1));
                    py = (int)(tile / (unsigned)tiles_x) * TILE_H + (int)(in / TILE_W);
                    if (px < p.width && py < p.height) {
                        have_job = true; pixel = (uint32_t)(py * p.width + px);
                        s = 0; bounce = 0; sum = mk3(0.0f, 0.0f, 0.0f);
                        start_path(p, px, py, pixel, sample_begin, ps); ++n_paths;
                    }                                        // else: off-image slot of an edge tile, fetch again next trip
                }
            }
        }
        if (__all_sync(GPT_FULL_MASK, done && !have_job)) break;
        // ---- arm and traverse the next ray of this lane's path (tracer.cu:72,94) ----
        if (have_job && bounce < p.depth) {
            ++n_rays;
            if (count_bounces && bounce < 16) atomicAdd(&sm.bounce_cnt[bounce], 1ull);
            trav_init(ts, stack, ps.o, ps.d, has_mesh);
            while (!trav_finished(ts)) {
                while (trav_wants_node(ts)) {
                    node_step(sc.nodes, ts, stack);
                    if (!__any_sync(__activemask(), ts.leaf >= 0)) break;
                }
                while (trav_wants_tri(ts)) tri_step(sc.tris, ts, stack, ps.o, ps.d, cull);
            }
        }
    }

    // ---- stats: one atomic per warp ----
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        n_rays += __shfl_xor_sync(GPT_FULL_MASK, n_rays, off);
        n_paths += __shfl_xor_sync(GPT_FULL_MASK, n_paths, off);
    }
    if (lane == 0) { atomicAdd(&stats[0], n_rays); atomicAdd(&stats[1], n_paths); }
    if (count_bounces) {
        __syncthreads();
        if (threadIdx.x < 16 && sm.bounce_cnt[threadIdx.x]) atomicAdd(&stats[2 + threadIdx.x], sm.bounce_cnt[threadIdx.x]);
    }
}

// ---- test hook: first bounce of every pixel (SURVEY 8-b gpt_get_primary_hits) ----
__global__ void __launch_bounds__(128)
k_primary(const DevScene sc, const RenderParams p, const unsigned long long sample,
          int* __restrict__ geom, int* __restrict__ tri_idx, int* __restrict__ prim_idx, float* __restrict__ t_out)
{
    __shared__ SmemScene sm;
    const float* s_spheres; const float* s_boxes;
    stage_scene(sc, sm, s_spheres, s_boxes);
    const int n = p.width * p.height;
    for (int pixel = blockIdx.x * blockDim.x + threadIdx.x; pixel < n; pixel += gridDim.x * blockDim.x) {
        const int px = pixel % p.width, py = pixel / p.width;
        PathState ps;
        start_path(p, px, py, (uint32_t)pixel, sample, ps);
        const SceneHit h = intersect_scene(sc, s_spheres, s_boxes, ps.o, ps.d, p.cull != 0);
        if (geom) geom[pixel] = h.geom;
        if (tri_idx) tri_idx[pixel] = h.geom == GEO_TRI ? __float_as_int(tri_row_a(sc.tris, h.prim).w) : -1;
        if (prim_idx) prim_idx[pixel] = (h.geom == GEO_SPHERE || h.geom == GEO_BOX) ? h.prim : -1;
        if (t_out) t_out[pixel] = h.t;
    }
}

// ---- test hook: nearest triangle hit of arbitrary rays ----
__global__ void __launch_bounds__(128)
k_trace_rays(const DevScene sc, const float* __restrict__ o, const float* __restrict__ d, const size_t n, const int cull,
             int* __restrict__ tri_idx, int* __restrict__ tri_addr, float* __restrict__ t_out)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        TriHit h; h.t = GPT_F32_MAX; h.tri = -1;
        if (sc.n_nodes > 0) {
            const f3 ro = mk3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), rd = mk3(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
            h = sc.wnodes ? traverse_bvh4(sc.wnodes, sc.tris, ro, rd, cull != 0) : traverse_bvh(sc.nodes, sc.tris, ro, rd, cull != 0);
        }
        if (tri_idx) tri_idx[i] = h.tri >= 0 ? __float_as_int(tri_row_a(sc.tris, h.tri).w) : -1;
        if (tri_addr) tri_addr[i] = h.tri >= 0 ? __float_as_int(__ldg(sc.tris + GPT_TRI_STRIDE * (size_t)h.tri + GPT_TRI_B).w) : -1;
        if (t_out) t_out[i] = h.t;
    }
}

// ---- resolve: mean = sum / count, clamp, pack (tracer.cu:390-398 applied once instead of per frame) ----
__global__ void __launch_bounds__(256)
k_resolve(const float4* __restrict__ accum, const int n, uint32_t* __restrict__ rgba8, float* __restrict__ rgb)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 a = accum[i];
        const float inv = a.w > 0.0f ? 1.0f / a.w : 0.0f;
        const float mx = a.x * inv, my = a.y * inv, mz = a.z * inv;
        if (rgb) { rgb[3 * i] = mx; rgb[3 * i + 1] = my; rgb[3 * i + 2] = mz; }
        if (rgba8) rgba8[i] = pack_rgb(clamp01(mx), clamp01(my), clamp01(mz));
    }
}

// ---- multi-GPU frame reduce fused with the running-sum update and the resolve ----
// Runs on the ROOT GPU.  frames[r] is rank r's per-frame float4 buffer: the root's own memory for r == root, a
// peer GPU's memory mapped through CUDA IPC otherwise, read with plain 16-byte loads over NVLink/NVSwitch.  Ranks
// are summed in rank order (deterministic, unlike a ring/tree reduce), added into the running sum, and the same
// thread resolves the pixel: one kernel replaces ncclReduce + accumulate + resolve (three passes over the frame).
#define GPT_MAX_RANKS 16
struct PeerFrames { const float4* f[GPT_MAX_RANKS]; int n; };

__global__ void __launch_bounds__(256)
k_reduce_resolve(const PeerFrames fr, const int n_pix, float4* __restrict__ total, uint32_t* __restrict__ rgba8, float* __restrict__ rgb)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_pix; i += gridDim.x * blockDim.x) {
        float4 v[GPT_MAX_RANKS];
#pragma unroll
        for (int r = 0; r < GPT_MAX_RANKS; ++r)
            if (r < fr.n) v[r] = __ldcs(fr.f[r] + i);          // all peer loads in flight before the first add
        float4 s = v[0];
#pragma unroll
        for (int r = 1; r < GPT_MAX_RANKS; ++r)
            if (r < fr.n) { s.x += v[r].x; s.y += v[r].y; s.z += v[r].z; s.w += v[r].w; }
        float4 a = s;
        if (total) { a = total[i]; a.x += s.x; a.y += s.y; a.z += s.z; a.w += s.w; total[i] = a; }
        const float inv = a.w > 0.0f ? 1.0f / a.w : 0.0f;
        const float mx = a.x * inv, my = a.y * inv, mz = a.z * inv;
        if (rgb) { rgb[3 * i] = mx; rgb[3 * i + 1] = my; rgb[3 * i + 2] = mz; }
        if (rgba8) rgba8[i] = pack_rgb(clamp01(mx), clamp01(my), clamp01(mz));
    }
}

cudaError_t gptk_reduce_resolve(const float4* const* frames, int n_frames, int n_pix, float4* total, uint32_t* rgba8, float* rgb,
                                int sm_count, cudaStream_t st)
{
    if (n_pix <= 0 || n_frames <= 0) return cudaSuccess;
    if (n_frames > GPT_MAX_RANKS) return cudaErrorInvalidValue;
    PeerFrames fr; fr.n = n_frames;
    for (int r = 0; r < GPT_MAX_RANKS; ++r) fr.f[r] = r < n_frames ? frames[r] : nullptr;
    int grid = (n_pix + 255) / 256; if (grid > sm_count * 8) grid = sm_count * 8;
    k_reduce_resolve<<<grid, 256, 0, st>>>(fr, n_pix, total, rgba8, rgb);
    return cudaGetLastError();
}

// ---------------------------------------------------------------- launchers
static int persistent_grid(const void* kernel, int threads, int sm_count)
{
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
    return sm_count * per_sm;     // a multiple of the SM count (148 on B200), every CTA resident
}

cudaError_t gptk_render_mega(const DevScene& sc, const RenderParams& p, float4* accum_f4, unsigned long long sample_begin,
                             uint32_t sample_count, unsigned int* work_counter, unsigned long long* stats,
                             int sm_count, cudaStream_t st)
{
    static int grid = 0;
    if (!grid) grid = persistent_grid((const void*)k_render_mega<false>, MEGA_THREADS, sm_count);
    cudaError_t e = cudaMemsetAsync(work_counter, 0, sizeof(unsigned int), st);
    if (e != cudaSuccess) return e;
    k_render_mega<false><<<grid, MEGA_THREADS, 0, st>>>(sc, p, accum_f4, nullptr, nullptr, 1ull, sample_begin, sample_count, work_counter, stats);
    return cudaGetLastError();
}

cudaError_t gptk_render_legacy(const DevScene& sc, const RenderParams& p, float* accum_v3, uint32_t* draw,
                               unsigned long long constant_pdf, unsigned int* work_counter, unsigned long long* stats,
                               int sm_count, cudaStream_t st)
{
    static int grid = 0;
    if (!grid) grid = persistent_grid((const void*)k_render_mega<true>, MEGA_THREADS, sm_count);
    cudaError_t e = cudaMemsetAsync(work_counter, 0, sizeof(unsigned int), st);
    if (e != cudaSuccess) return e;
    k_render_mega<true><<<grid, MEGA_THREADS, 0, st>>>(sc, p, nullptr, accum_v3, draw, constant_pdf, 0ull, 1u, work_counter, stats);
    return cudaGetLastError();
}

cudaError_t gptk_primary(const DevScene& sc, const RenderParams& p, unsigned long long sample,
                         int* geom, int* tri_idx, int* prim_idx, float* t, int sm_count, cudaStream_t st)
{
    const int n = p.width * p.height;
    int grid = (n + 127) / 128; if (grid > sm_count * 16) grid = sm_count * 16; if (grid < 1) grid = 1;
    k_primary<<<grid, 128, 0, st>>>(sc, p, sample, geom, tri_idx, prim_idx, t);
    return cudaGetLastError();
}

cudaError_t gptk_trace_rays(const DevScene& sc, const float* o, const float* d, size_t n, int cull,
                            int* tri_idx, int* tri_addr, float* t, int sm_count, cudaStream_t st)
{
    if (n == 0) return cudaSuccess;
    size_t g = (n + 127) / 128; if (g > (size_t)sm_count * 16) g = (size_t)sm_count * 16;
    k_trace_rays<<<(int)g, 128, 0, st>>>(sc, o, d, n, cull, tri_idx, tri_addr, t);
    return cudaGetLastError();
}

cudaError_t gptk_resolve(const float4* accum, int n, uint32_t* rgba8, float* rgb, int sm_count, cudaStream_t st)
{
    if (n <= 0) return cudaSuccess;
    int grid = (n + 255) / 256; if (grid > sm_count * 8) grid = sm_count * 8;
    k_resolve<<<grid, 256, 0, st>>>(accum, n, rgba8, rgb);
    return cudaGetLastError();
}
