// gpt_wavefront.cu -- wavefront form of the render pass (north_star item 2: "wavefront pipeline with compacted
// ray queues").  Same per-sample arithmetic as the megakernel (gpt_shade.cuh / gpt_traverse.cuh), so both
// produce bit-identical images; what changes is the schedule:
//
//   wave = a run of pixels x a run of global sample indices, up to 64 M paths   (path id = sample-major, then pixel)
//   bounce 0      : no generate pass -- the first extend and the first shade both derive the camera ray from the path id
//                   (tracer.cu:61-69; two Philox blocks cost less than writing and re-reading 52 B of state per path)
//   per bounce    : k_wf_extend_ww (default; k_wf_extend / k_wf_extend_pool are A/B schedules)
//                                persistent warps pull rays from the queue and traverse the BVH; a lane
//                                whose ray ends is re-armed with the next queued ray (a 32-byte load), so
//                                short rays never wait for long ones                  (cudaUtils.h:256-460)
//                   k_wf_shade   one thread per surviving path: sphere sweep, shading, next ray, and
//                                compaction of the survivors into the next queue      (tracer.cu:98-296)
//   k_wf_accumulate : the wave's finished samples, summed per pixel in sample order -> caller's accumulation buffer
//                     (tracer.cu:374,386-398); launches of several waves go through k_wf_collect + per-pixel launch sums
//
// ncu evidence for this choice over the megakernel is in profiles/ (r01_mega_*): with secondary rays the
// per-ray traversal length is heavy tailed (mean 10.8 node+triangle steps, ~41 for the slowest lane of a
// warp), and in a megakernel re-arming a finished lane costs a whole shading step.
#include "gpt_shade.cuh"
#include "gpt_kernels.h"
#include <cstdlib>
#include <cstring>
#include <cub/device/device_radix_sort.cuh>

#define WF_THREADS 128
// Path state (ray, mask, hit, queue entries) is touched ONCE per kernel, the scene again and again: with WF_STREAM the state goes
// through the streaming (evict-first) load / store forms so that it does not push tree nodes and triangles out of L1 / L2.
#ifndef WF_STREAM
#define WF_STREAM 0
#endif
// The ray of a path = ONE 32-byte record (origin xyz + mask x | direction xyz + mask y), read and written with one 256-bit
// access (LDG / STG .256, sm_100): path ids are scattered after the first bounce, so every access instruction costs a tag look-up
// per lane.
#ifndef WF_RAY256
#define WF_RAY256 1      // 0: the same record as two 128-bit accesses (A/B)
#endif
__device__ __forceinline__ void ld_ray32(const float4* __restrict__ ray, const uint32_t pid, float4& o, float4& d)
{
#if WF_RAY256
    asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(o.x), "=f"(o.y), "=f"(o.z), "=f"(o.w), "=f"(d.x), "=f"(d.y), "=f"(d.z), "=f"(d.w) : "l"(ray + 2 * (size_t)pid) : "memory");
#else
    o = ray[2 * (size_t)pid]; d = ray[2 * (size_t)pid + 1];
#endif
}
__device__ __forceinline__ void st_ray32(float4* __restrict__ ray, const uint32_t pid, const float4 o, const float4 d)
{
#if WF_RAY256
    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 :: "l"(ray + 2 * (size_t)pid), "f"(o.x), "f"(o.y), "f"(o.z), "f"(o.w), "f"(d.x), "f"(d.y), "f"(d.z), "f"(d.w) : "memory");
#else
    ray[2 * (size_t)pid] = o; ray[2 * (size_t)pid + 1] = d;
#endif
}
#if WF_STREAM
#define WF_LD(p) __ldcs(p)
#define WF_ST(p, v) __stcs((p), (v))
#else
#define WF_LD(p) (*(p))
#define WF_ST(p, v) (*(p) = (v))
#endif
#ifndef SHADE_MIN_BLOCKS
#define SHADE_MIN_BLOCKS 5
#endif
#ifndef WF_LEAF_SPLIT
#define WF_LEAF_SPLIT 0   // A/B: culled leaves in two passes (determinant filter, then full tests), see leaf_step_culled; measured 1.5 % slower
#endif
#ifndef WF_NODE_K
#define WF_NODE_K 3      // node-loop exit policy of k_wf_extend_ww, see there (profiles/r01_schedules.md #16)
#endif
#ifndef WF_MIN_BLOCKS
#define WF_MIN_BLOCKS 10   // 51-register cap, no spills; measured +1.5% over 8 (profiles/r01_sweep_extend.log)
#endif

struct SmemPrims {
    float4 posrad[GPT_MAX_SMEM_SPHERES];       // centre + radius of every sphere again, 16-byte aligned: the sweep reads ONE LDS.128 per sphere
    float spheres[GPT_MAX_SMEM_SPHERES * 11];
    float boxes[GPT_MAX_SMEM_BOXES * 13];
};

__device__ __forceinline__ void stage_prims(const DevScene& sc, SmemPrims& sm, const float*& s_spheres, const float*& s_boxes)
{
    s_spheres = sc.spheres; s_boxes = sc.boxes;
    if (sc.n_spheres <= GPT_MAX_SMEM_SPHERES) {
        for (int i = threadIdx.x; i < sc.n_spheres * 11; i += blockDim.x) sm.spheres[i] = sc.spheres[i];
        for (int i = threadIdx.x; i < sc.n_spheres; i += blockDim.x)
            sm.posrad[i] = make_float4(sc.spheres[11 * i], sc.spheres[11 * i + 1], sc.spheres[11 * i + 2], sc.spheres[11 * i + 3]);
        s_spheres = sm.spheres;
    }
    if (sc.n_boxes <= GPT_MAX_SMEM_BOXES) {
        for (int i = threadIdx.x; i < sc.n_boxes * 13; i += blockDim.x) sm.boxes[i] = sc.boxes[i];
        s_boxes = sm.boxes;
    }
    __syncthreads();
}

// counters layout (unsigned int): [0 .. depth]        n_q[b]   = paths alive at the start of bounce b  (b >= 1; bounce 0 = the whole wave)
//                                 [depth+1 .. 2depth+1] fetch[b] = extend's work-fetch cursor of bounce b
// path id = s_local * n_wave + i  <->  (global sample sample0 + s_local, pixel pix0 + i): sample-major, so the lanes of a warp
// hold neighbouring pixels of the same sample.  At bounce 0 there is no queue and no stored ray: path ids run 0 .. n_paths-1
// and the camera ray is a function of the id.
struct PrimaryGen { RenderParams p; unsigned long long sample0; int pix0; int n_wave; unsigned n_paths; };

__device__ __forceinline__ void primary_ray(const PrimaryGen& g, const uint32_t pid, PathState& ps)
{
    const int s_local = (int)(pid / (unsigned)g.n_wave);
    const int pixel = g.pix0 + (int)(pid - (unsigned)s_local * (unsigned)g.n_wave);
    start_path(g.p, pixel % g.p.width, pixel / g.p.width, (uint32_t)pixel, g.sample0 + (unsigned long long)s_local, ps);
}

// Persistent while-while traversal kernel (cudaUtils.h:256-460).  All lanes of a warp meet at the top of the outer loop once
// per "descend to the next leaf + test that leaf" round; idle lanes are re-armed there from the bounce's queue with one
// warp-aggregated atomic.  Five other schedules were built, checked bit-identical and measured slower -- round 1: voted single
// steps, pooled leaves per round (profiles/r01_schedules.md #6, #13); round 2: re-arming inside the node loop, and leaf phases
// pooled over the warp with one or three postponed leaves per lane (profiles/r02_schedules.md #43-#46: 25-60 % slower although
// a CPU replay of the warp schedule had predicted a gain) -- the code is in the history.
#ifndef WF_WIDE_MIN_BLOCKS
#define WF_WIDE_MIN_BLOCKS 10   // 48 registers, no spills; measured +2 % (2 M triangles) to +6 % (20 M) over 8 (profiles/r01_schedules.md #28)
#endif
// WIDE: the inner-node phase walks the 4-wide quantised tree (gpt_wide.cu, node4_step) instead of the binary one.
// PRIMARY: bounce 0 -- path ids come straight from the fetch cursor and the ray is the camera ray of that id (no stored ray yet).
template <bool CULL, bool WIDE, bool PRIMARY>      // back-face culling resolved at compile time: the flag is tested once per triangle otherwise
__global__ void __launch_bounds__(WF_THREADS, WIDE ? WF_WIDE_MIN_BLOCKS : WF_MIN_BLOCKS)
k_wf_extend_ww(const DevScene sc, const uint32_t* __restrict__ queue, const unsigned int* __restrict__ n_ptr,
               unsigned int* __restrict__ fetch, const float4* __restrict__ ray,
               int2* __restrict__ hit, const int refill_min, const int node_k, const PrimaryGen gen)
{
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    const unsigned n = PRIMARY ? gen.n_paths : *n_ptr;
    const bool do_cull = CULL;
    int stack[WIDE ? GPT_W4_STACK : GPT_STACK_SIZE];
    TravState ts;
    f3 o = mk3(0.f, 0.f, 0.f), d = mk3(0.f, 0.f, 1.f);
    uint32_t pid = 0;
    bool active = false, exhausted = false;
    ts.node = GPT_SENTINEL; ts.leaf = 0; ts.sp = 0; ts.t = GPT_F32_MAX; ts.tri = -1; ts.tri_cur = 0;

    while (true) {
        const unsigned m_idle = __ballot_sync(GPT_FULL_MASK, !active);
        if (m_idle != 0u && !exhausted && (__popc(m_idle) >= refill_min || m_idle == GPT_FULL_MASK)) {
            const unsigned k_idle = (unsigned)__popc(m_idle);
            unsigned base = 0;
            const int leader = __ffs(m_idle) - 1;
            if ((int)lane == leader) base = atomicAdd(fetch, k_idle);
            base = __shfl_sync(GPT_FULL_MASK, base, leader);
            if (base + k_idle >= n) exhausted = true;
            const unsigned j_lane = base + (unsigned)__popc(m_idle & lt_mask);
            if (!active) {
                const unsigned j = j_lane;
                if (j < n) {
                    if (PRIMARY) {
                        pid = j;
                        PathState ps; primary_ray(gen, pid, ps);
                        o = ps.o; d = ps.d;
                    } else {
                        // the ray records lie in QUEUE order (the shade kernel writes a survivor's record at its queue slot): no
                        // path id in between, and neighbouring lanes read neighbouring 32-byte records.  `queue` is only the
                        // permutation of the optional ray re-ordering (nullptr: slot = position)
                        pid = queue ? WF_LD(queue + j) : j;
                        float4 ro, rd;
                        ld_ray32(ray, pid, ro, rd);
                        o = mk3(ro.x, ro.y, ro.z); d = mk3(rd.x, rd.y, rd.z);
                    }
                    trav_init(ts, stack, o, d, true);
                    active = true;
                }
            }
        }
        if (__ballot_sync(GPT_FULL_MASK, active) == 0u) break;      // nothing armed: queue exhausted
        if (active) {
            while (trav_wants_node(ts)) {
                if (WIDE) node4_step(sc.wnodes, ts, stack); else node_step(sc.nodes, ts, stack);
                // reference policy (node_k = 1, cudaUtils.h:385-391): descend until every lane still in the loop holds a
                // leaf.  node_k = k leaves as soon as fewer than k lanes lack a leaf -- they resume after the leaf phase --
                // instead of making the whole warp wait for one or two stragglers (k = 2..3: -3 % extend time; k >= 8 loses)
                if (__popc(__ballot_sync(__activemask(), ts.leaf >= 0)) < node_k) break;
            }
            if (CULL && WF_LEAF_SPLIT) {
                if (trav_wants_tri(ts)) leaf_step_culled(sc.tris, ts, stack, o, d);     // one leaf per round, like the loop below
            } else {
                while (trav_wants_tri(ts)) {
                    tri_step(sc.tris, ts, stack, o, d, do_cull);
                    // a second leaf was postponed while this one waited: leave it for the next round (measured
                    // faster than draining it here: the round ends sooner for the rest of the warp)
                    if (ts.leaf < 0 && ts.tri_cur == ~ts.leaf) break;
                }
            }
            if (trav_finished(ts)) {
                WF_ST(hit + pid, make_int2(__float_as_int(ts.t), ts.tri));
                active = false;
            }
        }
    }
}

// One thread per path alive at bounce `bounce`: the rest of the bounce-loop body (tracer.cu:98-296), the path's radiance when
// it ends (tracer.cu:305 / :141, left in accu[pid]), and compaction of the survivors.
// FIRST: bounce 0 -- every path of the wave is alive, the ray is the camera ray of the path id (recomputed, not loaded), the
// mask is 1 (tracer.cu:68) and the accumulated colour starts at 0 (:69), so the slot is WRITTEN here, never read.
// FAST: GPT_FLAG_FAST shading arithmetic (gpt_math.cuh) -- opt-in, statistically equal, not bit-equal.
struct ShadeIo {
    const uint32_t* queue_in; unsigned int* counters; uint32_t* queue_out;
    const float4* ray_in; float4* ray_out; const float* mask_in; float* mask_out;      // records in queue order: 2 float4 + 1 float per slot
    const int2* hit; float4* accu;                                                       // hit: by the slot of the queue read; accu: by path id
};

#ifndef WF_SPHERE_LDS
#define WF_SPHERE_LDS 1
#endif
#ifndef WF_SHADE_PF
#define WF_SHADE_PF 1
#endif
template <int LEVEL> __device__ __forceinline__ void shade_prefetch(const void* p)
{
    if (LEVEL == 1) asm volatile("prefetch.global.L1 [%0];" :: "l"(p));
    if (LEVEL == 2) asm volatile("prefetch.global.L2 [%0];" :: "l"(p));
}

template <bool FIRST>
__device__ __forceinline__ void shade_load_path(const PrimaryGen& gen, const ShadeIo& io, const uint32_t pid, const unsigned slot, PathState& ps)
{
    if (FIRST) primary_ray(gen, pid, ps);
    else {
        float4 ro, rd;
        ld_ray32(io.ray_in, slot, ro, rd);
        ps.o = mk3(ro.x, ro.y, ro.z); ps.d = mk3(rd.x, rd.y, rd.z);
        ps.mask = mk3(ro.w, rd.w, WF_LD(io.mask_in + slot));
    }
    ps.accu = mk3(0.0f, 0.0f, 0.0f);      // accu stays in memory (shade_bounce, accu_slot)
}

// shading of one path + the survivor's state; returns whether it goes on to the next bounce
template <bool FIRST, bool FAST>
__device__ __forceinline__ bool shade_finish_path(const DevScene& sc, const PrimaryGen& gen, const ShadeIo& io, const float* s_spheres, const float* s_boxes,
                                                  const SceneHit h, const uint32_t pid, const uint32_t bounce, const bool last_bounce, PathState& ps)
{
    const uint32_t n_wave = (uint32_t)gen.n_wave;
    const uint32_t s_local = pid / n_wave;
    // accu[pid] is touched inside only on a miss or an emitting hit (always written at bounce 0); when the path is over its
    // radiance (tracer.cu:305, or bkColor on a miss :141) stays there until the wave's samples are summed per pixel
    const bool cont = shade_bounce<FIRST, FAST>(sc, gen.p, s_spheres, s_boxes, h, (uint32_t)gen.pix0 + (pid - s_local * n_wave), gen.sample0 + s_local, bounce, ps, io.accu + pid);
    return cont && !last_bounce;          // the survivor's record is stored by the caller, at the queue slot the compaction gives it
}

// survivors of a warp -> slots of the next bounce's queue: one atomic per warp.  Returns this lane's slot (only meaningful if it survives)
__device__ __forceinline__ unsigned shade_compact(const bool survive, unsigned int* n_out, const unsigned lane)
{
    const unsigned m = __ballot_sync(GPT_FULL_MASK, survive);
    unsigned base = 0;
    if (m) {
        const int leader = __ffs(m) - 1;
        if ((int)lane == leader) base = atomicAdd(n_out, (unsigned)__popc(m));
        base = __shfl_sync(GPT_FULL_MASK, base, leader);
    }
    return base + (unsigned)__popc(m & ((1u << lane) - 1u));
}

// the same for a whole CTA of 256 threads: ONE atomic per CTA and trip instead of one per warp.  All the queue atomics of a launch hit
// one address, and the L2's atomic unit serialises them (~0.85 cycles each: 1 M warp-level atomics of a 33 M-path launch are
// 0.45 ms of a 1.2 ms kernel, and every warp queues behind the others for its result).  Every thread of the CTA must call it.
__device__ __forceinline__ unsigned shade_compact_cta(const bool survive, unsigned int* n_out, const unsigned lane, unsigned* s_cnt, unsigned* s_base)
{
    const unsigned m = __ballot_sync(GPT_FULL_MASK, survive);
    const unsigned w = threadIdx.x >> 5;
    if (lane == 0) s_cnt[w] = (unsigned)__popc(m);
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned tot = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) { const unsigned t = s_cnt[k]; s_cnt[k] = tot; tot += t; }
        *s_base = tot ? atomicAdd(n_out, tot) : 0u;
    }
    __syncthreads();
    return *s_base + s_cnt[w] + (unsigned)__popc(m & ((1u << lane) - 1u));
}

#ifndef WF_SHADE_CTA_AGG
#define WF_SHADE_CTA_AGG 1
#endif
template <bool FIRST, bool FAST>
__global__ void __launch_bounds__(256, SHADE_MIN_BLOCKS)
k_wf_shade(const DevScene sc, const PrimaryGen gen, const uint32_t bounce, const ShadeIo io,
           unsigned long long* __restrict__ stats)
{
    __shared__ SmemPrims sm;
    const float* s_spheres; const float* s_boxes;
    stage_prims(sc, sm, s_spheres, s_boxes);
    const RenderParams& p = gen.p;
    const unsigned n = FIRST ? gen.n_paths : io.counters[bounce];
    unsigned int* n_out = io.counters + bounce + 1;
    const unsigned lane = threadIdx.x & 31u;
    const bool has_mesh = sc.n_nodes > 0;
    const bool last_bounce = bounce + 1u >= p.depth;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        atomicAdd(&stats[0], (unsigned long long)n);
        if (FIRST) atomicAdd(&stats[1], (unsigned long long)n);
        if ((p.flags & RP_COUNT_BOUNCES) && bounce < 16) atomicAdd(&stats[2 + bounce], (unsigned long long)n);
    }
    // whole warps iterate together so the ballots below are warp-uniform
    __shared__ unsigned s_cnt[8], s_base;
    const unsigned n_round = WF_SHADE_CTA_AGG ? ((n + 255u) & ~255u) : ((n + 31u) & ~31u);      // CTA-uniform trip counts for the barriers
    const unsigned stride = gridDim.x * blockDim.x;
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    // Every record of a trip is read at the trip's own queue position i (path id, ray, mask, hit: neighbouring lanes read neighbouring
    // records, nothing depends on a fetched id); WF_SHADE_PF = 1 / 2 asks L1 / L2 for the next trip's records while this one computes.
    for (; i < n_round; i += stride) {
        bool survive = false, valid = i < n;
        uint32_t pid = 0;
        PathState ps;
        SceneHit h; h.t = GPT_F32_MAX; h.geom = GEO_NONE; h.prim = -1;
        if (valid) {
            pid = FIRST ? i : WF_LD(io.queue_in + i);
            shade_load_path<FIRST>(gen, io, pid, i, ps);
            float tri_t = GPT_F32_MAX; int tri = -1;
            if (has_mesh) { const int2 hh = WF_LD(io.hit + i); tri_t = __int_as_float(hh.x); tri = hh.y; }
            if (WF_SHADE_PF && i + stride < n) {
                const unsigned i2 = i + stride;
                if (!FIRST) { shade_prefetch<WF_SHADE_PF>(io.ray_in + 2 * (size_t)i2); if ((lane & 7u) == 0u) { shade_prefetch<WF_SHADE_PF>(io.mask_in + i2); shade_prefetch<WF_SHADE_PF>(io.queue_in + i2); } }
                if (has_mesh && (lane & 3u) == 0u) shade_prefetch<WF_SHADE_PF>(io.hit + i2);
            }
            // spheres staged in shared memory: the sweep reads them through the __shared__ array itself (LDS.128), not through the
            // generic pointer, whose 4 x 32-bit LD per sphere wait on the long scoreboard (ncu: the kernel's largest stall)
            if (WF_SPHERE_LDS && sc.n_spheres <= GPT_MAX_SMEM_SPHERES) h = finish_scene_hit_smem<FAST>(sc, sm.posrad, s_boxes, tri_t, tri, ps.o, ps.d);
            else h = finish_scene_hit<FAST>(sc, s_spheres, s_boxes, tri_t, tri, ps.o, ps.d);
        }
        if (valid) survive = shade_finish_path<FIRST, FAST>(sc, gen, io, s_spheres, s_boxes, h, pid, bounce, last_bounce, ps);
        const unsigned slot = WF_SHADE_CTA_AGG ? shade_compact_cta(survive, n_out, lane, s_cnt, &s_base) : shade_compact(survive, n_out, lane);
        if (survive) {
            st_ray32(io.ray_out, slot, make_float4(ps.o.x, ps.o.y, ps.o.z, ps.mask.x), make_float4(ps.d.x, ps.d.y, ps.d.z, ps.mask.y));
            WF_ST(io.mask_out + slot, ps.mask.z);
            WF_ST(io.queue_out + slot, pid);
        }
    }
}

// Launches of SEVERAL waves: adds the finished paths of a wave to the per-pixel launch sums, samples in increasing global
// index: the same left-to-right float sum the megakernel keeps in a register, so both pipelines round identically.
// The last block to pass also re-zeroes the wave's queue counters for the next wave.
__global__ void __launch_bounds__(256)
k_wf_collect(const float4* __restrict__ accu, const int n_samples, const int pix0, const int n_wave, float4* __restrict__ lsum,
             unsigned int* __restrict__ counters, const int n_counters)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_wave; i += gridDim.x * blockDim.x) {
        float4 s = lsum[pix0 + i];
        for (int k = 0; k < n_samples; ++k) {
            const float4 a = accu[(size_t)k * n_wave + i];
            s.x += a.x; s.y += a.y; s.z += a.z;
        }
        lsum[pix0 + i] = s;
    }
    if (blockIdx.x == 0) for (int k = threadIdx.x; k < n_counters; k += blockDim.x) counters[k] = 0u;     // every shade kernel of the wave has finished
}

// Per-pixel sums of the launch -> caller's buffer.  LEGACY: running mean + clamp + pack (tracer.cu:386-398).
// DIRECT (the launch was ONE wave covering the frame, the usual case): the pixel's samples are summed here, in increasing
// global index, straight from the wave's accu[] -- no launch-sum buffer, no collect pass (tracer.cu:374).  Otherwise src = the
// launch sums k_wf_collect kept.  depth == 0: every path returns accucolor = 0 (tracer.cu:72,305): the launcher takes the
// launch-sum form with a zeroed buffer.
template <bool LEGACY, bool DIRECT>
__global__ void __launch_bounds__(256)
k_wf_accumulate(const float4* __restrict__ src, const int n_pix, const uint32_t sample_count, float4* __restrict__ accum_f4,
                float* __restrict__ accum_v3, uint32_t* __restrict__ draw, const unsigned long long constant_pdf,
                unsigned int* __restrict__ counters, const int n_counters)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_pix; i += gridDim.x * blockDim.x) {
        float4 s;
        if (DIRECT) {
            s = make_float4(0.0f, 0.0f, 0.0f, 0.0f);                // starts at +0 like the zeroed launch-sum buffer: same roundings
            for (uint32_t k = 0; k < sample_count; ++k) {
                const float4 a = src[(size_t)k * n_pix + i];
                s.x += a.x; s.y += a.y; s.z += a.z;
            }
        } else {
            s = src[i];
        }
        if (LEGACY) {
            const float nm1 = (float)(constant_pdf - 1ull), inv = 1.0f / (float)constant_pdf;
            float* a = accum_v3 + 3 * (size_t)i;
            const float ax = fminf(fmaxf((a[0] * nm1 + s.x) * inv, 0.0f), 1.0f);
            const float ay = fminf(fmaxf((a[1] * nm1 + s.y) * inv, 0.0f), 1.0f);
            const float az = fminf(fmaxf((a[2] * nm1 + s.z) * inv, 0.0f), 1.0f);
            a[0] = ax; a[1] = ay; a[2] = az;
            if (draw) {
                const unsigned char cx = (unsigned char)(int)(255.0f * ax), cy = (unsigned char)(int)(255.0f * ay), cz = (unsigned char)(int)(255.0f * az);
                draw[i] = ((uint32_t)cz << 16) | ((uint32_t)cy << 8) | (uint32_t)cx;
            }
        } else {
            float4 a = accum_f4[i];
            a.x += s.x; a.y += s.y; a.z += s.z; a.w += (float)sample_count;
            accum_f4[i] = a;
        }
    }
    if (DIRECT && blockIdx.x == 0) for (int k = threadIdx.x; k < n_counters; k += blockDim.x) counters[k] = 0u;
}

// ---------------------------------------------------------------- launcher
static int wf_knob(const char* name, int dflt)
{
    const char* e = getenv(name);
    int v = e ? atoi(e) : dflt;
    return v < 1 ? 1 : (v > 32 ? 32 : v);
}

size_t gptk_wave_counter_count(unsigned int depth) { return 2 * ((size_t)depth + 1); }

// ---- optional ray reordering (GPT_WF_SORT): secondary rays are incoherent; sorting the queue by a key made of the
// origin's cell (7 bits per axis of the mesh bounds, Morton order) and the direction's octant puts rays that walk the
// same part of the tree into the same warp.  Only the ORDER in which the extend kernel fetches rays changes: hit[pid]
// is per ray, so images are bit-identical.  The count of live rays stays on the device; entries past it get the
// largest key and are never fetched.
#define WF_SORT_BITS 30     // upper bound (temporary storage is sized for it); the bits actually sorted depend on the mode
__device__ __forceinline__ unsigned wf_spread10(unsigned v)      // 10 bits -> every third bit
{
    v &= 0x3ffu; v = (v | (v << 16)) & 0x030000ffu; v = (v | (v << 8)) & 0x0300f00fu; v = (v | (v << 4)) & 0x030c30c3u; v = (v | (v << 2)) & 0x09249249u;
    return v;
}
__host__ __device__ __forceinline__ int wf_sort_bits(const int mode) { return mode >= 8 ? 16 : mode >= 3 ? 30 : 24; }
// key = Morton cell of the origin (ob bits per axis of the mesh bounds) above Morton cell of the direction (db bits per axis):
// mode 1: 7 + 1 (octant) = 24 bits   mode 2: 6 + 2 = 24   mode 3: 7 + 3 = 30   mode 4: 9 + 1 = 30   mode 5: 8 + 2 = 30   mode 6: 6 + 4 = 30   mode 7: 5 + 5 = 30   mode 8: 4 + 1 = 15 (two radix passes)   mode 9: 3 + 2 = 15
__global__ void __launch_bounds__(256)
k_wf_sort_keys(const DevScene sc, uint32_t* __restrict__ iota, const unsigned int* __restrict__ n_ptr, const unsigned n_slots,
               const float4* __restrict__ ray, uint32_t* __restrict__ keys, const int mode)
{
    const unsigned n = *n_ptr;
    const unsigned kmax = 0xffffffffu >> (32 - wf_sort_bits(mode));
    const int ob = mode == 1 ? 7 : mode == 2 ? 6 : mode == 3 ? 7 : mode == 4 ? 9 : mode == 5 ? 8 : mode == 6 ? 6 : mode == 7 ? 5 : mode == 8 ? 4 : 3;
    const int db = mode == 1 ? 1 : mode == 2 ? 2 : mode == 3 ? 3 : mode == 4 ? 1 : mode == 5 ? 2 : mode == 6 ? 4 : mode == 7 ? 5 : mode == 8 ? 1 : 2;
    // mesh bounds = union of the root's two child boxes
    const float4 n0 = __ldg(sc.nodes), n1 = __ldg(sc.nodes + 1), nz = __ldg(sc.nodes + 2);
    const float lox = fminf(n0.x, n1.x), hix = fmaxf(n0.y, n1.y), loy = fminf(n0.z, n1.z), hiy = fmaxf(n0.w, n1.w);
    const float loz = fminf(nz.x, nz.z), hiz = fmaxf(nz.y, nz.w);
    const float oc = (float)(1 << ob), omax = oc - 1.0f;
    const float sx = oc * 0.9999f / fmaxf(hix - lox, 1e-20f), sy = oc * 0.9999f / fmaxf(hiy - loy, 1e-20f), sz = oc * 0.9999f / fmaxf(hiz - loz, 1e-20f);
    const float ds = (float)(1 << db) * 0.4999f, dmax = (float)((1 << db) - 1);
    for (unsigned j = blockIdx.x * blockDim.x + threadIdx.x; j < n_slots; j += gridDim.x * blockDim.x) {
        unsigned key = kmax;
        iota[j] = j;                                  // the values the keys are sorted with: queue slots
        if (j < n) {
            float4 o, d;
            ld_ray32(ray, j, o, d);
            const unsigned cx = (unsigned)fminf(fmaxf((o.x - lox) * sx, 0.0f), omax);
            const unsigned cy = (unsigned)fminf(fmaxf((o.y - loy) * sy, 0.0f), omax);
            const unsigned cz = (unsigned)fminf(fmaxf((o.z - loz) * sz, 0.0f), omax);
            const unsigned qx = (unsigned)fminf(fmaxf((d.x + 1.0f) * ds, 0.0f), dmax);
            const unsigned qy = (unsigned)fminf(fmaxf((d.y + 1.0f) * ds, 0.0f), dmax);
            const unsigned qz = (unsigned)fminf(fmaxf((d.z + 1.0f) * ds, 0.0f), dmax);
            const unsigned om = wf_spread10(cx) | (wf_spread10(cy) << 1) | (wf_spread10(cz) << 2);
            const unsigned dm = wf_spread10(qx) | (wf_spread10(qy) << 1) | (wf_spread10(qz) << 2);
            key = (om << (3 * db)) | dm;
            if (key >= kmax) key = kmax - 1u;
        }
        keys[j] = key;
    }
}
size_t gptk_wave_sort_tmp_bytes(size_t capacity)
{
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                    (long long)capacity, 0, WF_SORT_BITS, (cudaStream_t)0);
    return bytes;
}

// launch configuration of one context (grids follow from the device's SM count and the kernels' occupancy; the knobs are read
// from the environment once per context)
void gptk_wave_launch_cfg(WaveLaunchCfg* cfg, int sm_count)
{
    auto per_sm = [](const void* k, int threads, int dflt) { int v = 0; return (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, k, threads, 0) == cudaSuccess && v >= 1) ? v : dflt; };
    cfg->sm_count = sm_count;
    cfg->ext_grid = sm_count * per_sm((const void*)k_wf_extend_ww<true, false, false>, WF_THREADS, 1);
    cfg->wide_grid = sm_count * per_sm((const void*)k_wf_extend_ww<true, true, false>, WF_THREADS, 1);
    // shade grid: whole waves of resident CTAs (occupancy of the kernel as built x GPT_SHADE_WAVES), so no partial last wave
    const char* e = getenv("GPT_SHADE_WAVES");
    const int waves = e ? atoi(e) : 16;      // measured 1 / 2 / 4 / 8 / 16 / 32 / 64 / one trip per thread: 11.46 / 11.27 / 11.0 / 10.82 / 10.75 / 10.87 / 11.09 / 11.92 ms
    cfg->shade_grid = waves <= 0 ? 0x7fffffff : sm_count * per_sm((const void*)k_wf_shade<false, false>, 256, 4) * (waves > 4096 ? 4096 : waves);
    const int m = wf_knob("GPT_WF_SORT", 3);
    cfg->sort_mode = m > 9 ? 9 : m;
    cfg->refill_min = wf_knob("GPT_WF_REFILL_MIN", 8);
    cfg->node_k_env = getenv("GPT_WF_NODE_K") ? wf_knob("GPT_WF_NODE_K", WF_NODE_K) : 0;
    const char* w = getenv("GPT_WF_WAVE_PATHS");
    const long long v = w ? atoll(w) : 0;
    cfg->wave_cap_env = (size_t)(v < 0 ? 0 : v);
    cfg->ready = 1;
}

template <bool PRIMARY>
static void launch_extend(const DevScene& sc, const RenderParams& p, const WaveLaunchCfg& cfg, const uint32_t* q, unsigned int* n_ptr, unsigned int* fetch,
                          const float4* ray_in, const WaveBuffers& wb, int node_k, const PrimaryGen& gen, cudaStream_t st)
{
    if (sc.wnodes) {
        if (p.cull) k_wf_extend_ww<true, true, PRIMARY><<<cfg.wide_grid, WF_THREADS, 0, st>>>(sc, q, n_ptr, fetch, ray_in, wb.hit, cfg.refill_min, node_k, gen);
        else k_wf_extend_ww<false, true, PRIMARY><<<cfg.wide_grid, WF_THREADS, 0, st>>>(sc, q, n_ptr, fetch, ray_in, wb.hit, cfg.refill_min, node_k, gen);
    } else {
        if (p.cull) k_wf_extend_ww<true, false, PRIMARY><<<cfg.ext_grid, WF_THREADS, 0, st>>>(sc, q, n_ptr, fetch, ray_in, wb.hit, cfg.refill_min, node_k, gen);
        else k_wf_extend_ww<false, false, PRIMARY><<<cfg.ext_grid, WF_THREADS, 0, st>>>(sc, q, n_ptr, fetch, ray_in, wb.hit, cfg.refill_min, node_k, gen);
    }
}

// wb.counters must be all zero on entry; they are all zero again when the enqueued work has run (the last kernel of every wave
// re-zeroes them), so a call costs 2 x depth + 1 launches when one wave covers the frame.
cudaError_t gptk_render_wavefront(const DevScene& sc, const RenderParams& p, const WaveBuffers& wb, const WaveLaunchCfg& cfg,
                                  float4* accum_f4, float* accum_v3, uint32_t* draw, unsigned long long constant_pdf,
                                  unsigned long long sample_begin, uint32_t sample_count,
                                  unsigned long long* stats, cudaStream_t st, unsigned long long* launches,
                                  KernelTimer* kt)
{
    const int sm_count = cfg.sm_count;
    // node-loop exit policy of k_wf_extend_ww (see there).  Small scenes spend half their steps in leaves and want full leaf
    // phases (k = 3); larger scenes make 15-40 node visits per 1-2 triangle tests, so lanes that hold a leaf should not wait
    // for the rest of the warp to find one.  Measured extend time, k = 3 / 6 / 10 / 16 (profiles/r01_schedules.md #39): bench scene
    // (5 k triangles) 16.5 / - / 17.2 (k = 8) / 19.7; dragon 100 k 16.3 / 15.6 / 15.2 / 15.4; 16 cars 157 k 41.1 / 38.3 / 37.7 / 39.4;
    // dragon 1.6 M 23.3 / 21.3 / 20.7 / 20.9; synthetic 2 M 20.8 / 19.2 / 18.4 / 18.1; 20 M 37.6 / 35.2 / 33.9 / 33.2
    // Small scenes on a tree with small leaves (the device builder's, ~1.8 triangle references per inner node against ~3 for the
    // reference's SBVH) sit in between: bench scene on the rebuilt tree, k = 2 / 3 / 4 / 6 / 8: 13.03 / 12.54 / 12.36 / 12.34 / 12.52 ms
    // (profiles/r02_schedules.md #58), while the SBVH of the same mesh is best at 3
    const int small_k = (2 * (long long)sc.n_tris <= 5 * (long long)sc.n_nodes) ? 5 : WF_NODE_K;
    const int node_k = cfg.node_k_env ? cfg.node_k_env : (sc.n_tris < (64 << 10) ? small_k : (sc.n_tris < (2 << 20) ? 8 : 16));      // 8: between the reference-built dragon's best (6) and the device-built trees' (10)
    const int n_pix = p.width * p.height;
    const int n_counters = (int)gptk_wave_counter_count(p.depth);
    // A wave = a run of consecutive pixels x a run of consecutive samples, as many paths as the scratch arrays
    // hold (104 B/path).  Big waves matter: every launch has a fixed cost (launch gap + the tail where the
    // persistent extend kernel drains), measured at ~19 us, and a 1080p x 1 spp wave needs 17 launches.
    size_t cap = wb.capacity;
    if (cfg.wave_cap_env && cfg.wave_cap_env < cap) cap = cfg.wave_cap_env < 1024 ? 1024 : cfg.wave_cap_env;
    const int wave_pix = (int)((size_t)n_pix < cap ? (size_t)n_pix : cap);
    const uint32_t wave_spp_max = (uint32_t)(cap / (size_t)wave_pix);      // >= 1
    int agrid = (n_pix + 255) / 256; if (agrid > sm_count * 8) agrid = sm_count * 8; if (agrid < 1) agrid = 1;
    const bool legacy = accum_v3 != nullptr;
    const bool has_mesh = sc.n_nodes > 0;
    // one wave covers the launch: its samples are summed straight from accu[] by the accumulate kernel
    const bool direct = p.depth > 0 && wave_pix == n_pix && sample_count <= wave_spp_max;
    cudaError_t e;
    if (!direct && (e = cudaMemsetAsync(wb.lsum, 0, sizeof(float4) * (size_t)n_pix, st)) != cudaSuccess) return e;
    for (int pix0 = 0; pix0 < n_pix && p.depth > 0; pix0 += wave_pix) {
        const int n_wave = n_pix - pix0 < wave_pix ? n_pix - pix0 : wave_pix;
        for (uint32_t s0 = 0; s0 < sample_count; s0 += wave_spp_max) {
            const uint32_t ns = sample_count - s0 < wave_spp_max ? sample_count - s0 : wave_spp_max;
            const size_t n_paths = (size_t)ns * n_wave;
            int grid = (int)((n_paths + 255) / 256); if (grid > sm_count * 8) grid = sm_count * 8; if (grid < 1) grid = 1;
            int sgrid = (int)((n_paths + 255) / 256); if (sgrid > cfg.shade_grid) sgrid = cfg.shade_grid; if (sgrid < 1) sgrid = 1;
            int cgrid = (n_wave + 255) / 256; if (cgrid > sm_count * 8) cgrid = sm_count * 8; if (cgrid < 1) cgrid = 1;
            PrimaryGen gen; gen.p = p; gen.sample0 = sample_begin + s0; gen.pix0 = pix0; gen.n_wave = n_wave; gen.n_paths = (unsigned)n_paths;
            for (uint32_t b = 0; b < p.depth; ++b) {
                // path ids, rays and masks of the paths alive at bounce b lie in queue order in the buffers of parity b & 1; the shade
                // kernel of bounce b writes the survivors' records into the buffers of the other parity
                const uint32_t* q_in = b == 0 ? nullptr : wb.queue[b & 1u];
                const float4* ray_in = wb.ray + (size_t)(b & 1u) * 2 * wb.capacity;
                float4* ray_out = wb.ray + (size_t)((b + 1u) & 1u) * 2 * wb.capacity;
                const float* mask_in = wb.mask_z + (size_t)(b & 1u) * wb.capacity;
                float* mask_out = wb.mask_z + (size_t)((b + 1u) & 1u) * wb.capacity;
                const uint32_t* q_ext = nullptr;    // order in which the extend kernel fetches the slots: queue order, or the re-ordering's permutation
                if (has_mesh && b > 0 && wb.sort_vals) {
                    if (kt) kt->tick(KT_OTHER, st);
                    uint32_t* iota = wb.queue[(b + 1u) & 1u];      // free until this bounce's shade kernel fills it
                    k_wf_sort_keys<<<grid, 256, 0, st>>>(sc, iota, wb.counters + b, (unsigned)n_paths, ray_in, wb.sort_keys[0], cfg.sort_mode);
                    size_t tb = wb.sort_tmp_bytes;
                    if ((e = cub::DeviceRadixSort::SortPairs(wb.sort_tmp, tb, (const uint32_t*)wb.sort_keys[0], wb.sort_keys[1], (const uint32_t*)iota, wb.sort_vals,
                                                             (long long)n_paths, 0, wf_sort_bits(cfg.sort_mode), st)) != cudaSuccess) return e;
                    *launches += 4;
                    q_ext = wb.sort_vals;
                }
                if (has_mesh) {
                    if (kt) kt->tick(KT_EXTEND, st);
                    if (b == 0) launch_extend<true>(sc, p, cfg, nullptr, wb.counters, wb.counters + p.depth + 1, ray_in, wb, node_k, gen, st);
                    else launch_extend<false>(sc, p, cfg, q_ext, wb.counters + b, wb.counters + p.depth + 1 + b, ray_in, wb, node_k, gen, st);
                    *launches += 1;
                }
                if (kt) kt->tick(KT_SHADE, st);
                ShadeIo io;
                io.queue_in = q_in; io.counters = wb.counters; io.queue_out = wb.queue[(b + 1u) & 1u];
                io.ray_in = ray_in; io.ray_out = ray_out; io.mask_in = mask_in; io.mask_out = mask_out; io.hit = wb.hit; io.accu = wb.accu;
                if (p.flags & RP_FAST) {
                    if (b == 0) k_wf_shade<true, true><<<sgrid, 256, 0, st>>>(sc, gen, b, io, stats);
                    else k_wf_shade<false, true><<<sgrid, 256, 0, st>>>(sc, gen, b, io, stats);
                } else {
                    if (b == 0) k_wf_shade<true, false><<<sgrid, 256, 0, st>>>(sc, gen, b, io, stats);
                    else k_wf_shade<false, false><<<sgrid, 256, 0, st>>>(sc, gen, b, io, stats);
                }
                *launches += 1;
            }
            if (!direct) {
                if (kt) kt->tick(KT_COLLECT, st);
                k_wf_collect<<<cgrid, 256, 0, st>>>(wb.accu, (int)ns, pix0, n_wave, wb.lsum, wb.counters, n_counters);
                *launches += 1;
            }
            if ((e = cudaGetLastError()) != cudaSuccess) return e;
        }
    }
    if (kt) kt->tick(KT_ACCUMULATE, st);
    if (direct) {
        if (legacy) k_wf_accumulate<true, true><<<agrid, 256, 0, st>>>(wb.accu, n_pix, sample_count, nullptr, accum_v3, draw, constant_pdf, wb.counters, n_counters);
        else k_wf_accumulate<false, true><<<agrid, 256, 0, st>>>(wb.accu, n_pix, sample_count, accum_f4, nullptr, nullptr, 1ull, wb.counters, n_counters);
    } else {
        if (legacy) k_wf_accumulate<true, false><<<agrid, 256, 0, st>>>(wb.lsum, n_pix, sample_count, nullptr, accum_v3, draw, constant_pdf, wb.counters, n_counters);
        else k_wf_accumulate<false, false><<<agrid, 256, 0, st>>>(wb.lsum, n_pix, sample_count, accum_f4, nullptr, nullptr, 1ull, wb.counters, n_counters);
    }
    *launches += 1;
    if (kt) kt->tick(KT_END, st);
    return cudaGetLastError();
}
