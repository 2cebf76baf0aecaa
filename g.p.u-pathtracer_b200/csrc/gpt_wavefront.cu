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
    float spheres[GPT_MAX_SMEM_SPHERES * 11];
    float boxes[GPT_MAX_SMEM_BOXES * 13];
};

__device__ __forceinline__ void stage_prims(const DevScene& sc, SmemPrims& sm, const float*& s_spheres, const float*& s_boxes)
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
// warp-aggregated atomic.  Two other schedules were built and measured slower in round 1 (voted single steps, pooled leaves:
// profiles/r01_schedules.md #6, #13; the code is in the history).
#ifndef WF_WIDE_MIN_BLOCKS
#define WF_WIDE_MIN_BLOCKS 10   // 48 registers, no spills; measured +2 % (2 M triangles) to +6 % (20 M) over 8 (profiles/r01_schedules.md #28)
#endif
// WIDE: the inner-node phase walks the 4-wide quantised tree (gpt_wide.cu, node4_step) instead of the binary one.
// PRIMARY: bounce 0 -- path ids come straight from the fetch cursor and the ray is the camera ray of that id (no stored ray yet).
template <bool CULL, bool WIDE, bool PRIMARY>      // back-face culling resolved at compile time: the flag is tested once per triangle otherwise
__global__ void __launch_bounds__(WF_THREADS, WIDE ? WF_WIDE_MIN_BLOCKS : WF_MIN_BLOCKS)
k_wf_extend_ww(const DevScene sc, const uint32_t* __restrict__ queue, const unsigned int* __restrict__ n_ptr,
               unsigned int* __restrict__ fetch, const float4* __restrict__ ray_o, const float4* __restrict__ ray_d,
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
            unsigned base = 0;
            const int leader = __ffs(m_idle) - 1;
            if ((int)lane == leader) base = atomicAdd(fetch, (unsigned)__popc(m_idle));
            base = __shfl_sync(GPT_FULL_MASK, base, leader);
            if (base + (unsigned)__popc(m_idle) >= n) exhausted = true;
            if (!active) {
                const unsigned j = base + (unsigned)__popc(m_idle & lt_mask);
                if (j < n) {
                    if (PRIMARY) {
                        pid = j;
                        PathState ps; primary_ray(gen, pid, ps);
                        o = ps.o; d = ps.d;
                    } else {
                        pid = queue[j];
                        const float4 ro = ray_o[pid], rd = ray_d[pid];
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
                    if (ts.leaf < 0 && ts.tri_cur == GPT_LEAF_FIRST(~ts.leaf)) break;
                }
            }
            if (trav_finished(ts)) {
                hit[pid] = make_int2(__float_as_int(ts.t), ts.tri);
                active = false;
            }
        }
    }
}

// ---- pooled schedule (default from round 2; GPT_WF_EXTEND=ww selects the kernel above) -----------------------------------
// Why: in k_wf_extend_ww every lane walks its OWN leaf, so a leaf phase costs max-leaf-size trips while the lanes without a leaf,
// or with a short one, idle, and a lane that has met its second leaf idles through the rest of the node phase.  Measured (ncu,
// round 1): 17 of 32 lanes per instruction on the 5 k-triangle bench scene, 8-11 on multi-million-triangle scenes; a CPU replay
// of the warp schedule on real secondary rays (profiles/tools/warp_sim.cpp, profiles/r02_schedules.md) put the triangle phases
// at 5-13 lanes and pointed at this schedule (-15 % warp instructions per ray on the bench scene, -30 % from 100 k triangles up):
//   * a lane POSTPONES up to three leaves (+ the one it stalls on) and keeps descending: the node phase keeps its lanes;
//   * finished lanes are re-armed from the queue INSIDE the node loop, as soon as `refill_min` of them idle;
//   * the leaf phase starts when `hold_k` lanes hold leaves (or nobody can descend) and is POOLED: the warp's pending
//     (ray, triangle) pairs are laid end to end in shared memory (leaf sizes ride in the leaf pointers, so nothing is fetched for
//     that) and handed out 32 at a time, every lane testing one pair with the owner's ray read from shared memory.
// Same nearest hit, same tie-breaking as the reference loop (cudaUtils.h:400-446): a lane's leaves are tested in the order it met
// them and the winner is the lexicographic minimum of (t, position in that order) over the pairs with 0 < t < hitT -- what the
// sequential strict `t < hitT` update yields.  Postponing more leaves only delays the shrinking of hitT: subtrees the reference
// culls may be visited, but everything in them lies behind the final hit (their entry distance exceeded an earlier hitT).
#ifndef WF_PQ_MIN_BLOCKS
#define WF_PQ_MIN_BLOCKS 10
#endif
#ifndef WF_DEFAULT_SCHEDULE
#define WF_DEFAULT_SCHEDULE 0     // 0 = ww (measured fastest so far), 3 / 1 = pooled with 3 / 1 postponed leaves, 2 = in-loop re-arming
#endif
#define WF_PQ_CAP 256            // pairs laid out per chunk and warp
#define WF_PQ_MAXQ 3             // leaves a lane may postpone (the one it stalls on comes on top)
struct PqShared {
    float4 ray[WF_THREADS / 32][32][2];              // o.xyz + hitT, d.xyz + unused: the lane's ray, written when it is armed
    int leafq[WF_THREADS / 32][WF_PQ_MAXQ + 1][32];  // postponed leaves (x = ~pointer) per lane, oldest first
    unsigned item[WF_THREADS / 32][WF_PQ_CAP];       // owner lane << 27 | packed triangle
    unsigned long long best[WF_THREADS / 32][32];    // bits(t) << 32 | position of the pair: min = nearest, first met
};

__device__ __forceinline__ int pq_leaf_count(const float4* __restrict__ tris, const int x)
{
    const int f = (int)GPT_LEAF_SIZE_FIELD(x);
    return f < 15 ? f + 1 : __float_as_int(__ldg(tris + 3 * GPT_LEAF_FIRST(x) + 2).w) + 1;      // 16 or more: the first triangle knows
}

// MQ: leaves a lane postpones before it stalls (1 = the reference's single postponed leaf, cudaUtils.h:362-369).
// POOLED: leaf phase pooled over the warp (above) or every lane walking its own leaves (the schedule of k_wf_extend_ww with the
// re-arming moved into the node loop).
template <bool CULL, bool WIDE, bool PRIMARY, int MQ, bool POOLED>
__global__ void __launch_bounds__(WF_THREADS, WF_PQ_MIN_BLOCKS)
k_wf_extend_pq(const DevScene sc, const uint32_t* __restrict__ queue, const unsigned int* __restrict__ n_ptr,
               unsigned int* __restrict__ fetch, const float4* __restrict__ ray_o, const float4* __restrict__ ray_d,
               int2* __restrict__ hit, const int refill_min, const int hold_k, const PrimaryGen gen)
{
    __shared__ PqShared sm;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    const unsigned n = PRIMARY ? gen.n_paths : *n_ptr;
    float4* my_ray = sm.ray[warp][lane];
    int* my_q = &sm.leafq[warp][0][lane];           // entry k at my_q[32 * k]
    unsigned* items = sm.item[warp];
    int stack[WIDE ? GPT_W4_STACK : GPT_STACK_SIZE];
    TravState ts;
    int nq = 0;
    uint32_t pid = 0;
    bool active = false, exhausted = false;
    ts.node = GPT_SENTINEL; ts.leaf = 0; ts.sp = 0; ts.t = GPT_F32_MAX; ts.tri = -1; ts.tri_cur = 0;
    ts.idx = ts.idy = ts.idz = 1.0f; ts.oodx = ts.oody = ts.oodz = 0.0f;

    while (true) {
        // ------------------------------------------------------------ node phase
        while (true) {
            const unsigned m_idle = __ballot_sync(GPT_FULL_MASK, !active);
            if (m_idle != 0u && !exhausted && (__popc(m_idle) >= refill_min || m_idle == GPT_FULL_MASK)) {
                unsigned base = 0;
                const int leader = __ffs(m_idle) - 1;
                if ((int)lane == leader) base = atomicAdd(fetch, (unsigned)__popc(m_idle));
                base = __shfl_sync(GPT_FULL_MASK, base, leader);
                if (base + (unsigned)__popc(m_idle) >= n) exhausted = true;
                if (!active) {
                    const unsigned j = base + (unsigned)__popc(m_idle & lt_mask);
                    if (j < n) {
                        f3 o, d;
                        if (PRIMARY) {
                            pid = j;
                            PathState ps; primary_ray(gen, pid, ps);
                            o = ps.o; d = ps.d;
                        } else {
                            pid = queue[j];
                            const float4 ro = ray_o[pid], rd = ray_d[pid];
                            o = mk3(ro.x, ro.y, ro.z); d = mk3(rd.x, rd.y, rd.z);
                        }
                        trav_init(ts, stack, o, d, true);
                        my_ray[0] = make_float4(o.x, o.y, o.z, GPT_F32_MAX);
                        my_ray[1] = make_float4(d.x, d.y, d.z, 0.0f);
                        nq = 0;
                        active = true;
                    }
                }
            }
            // the leaf met in the previous trip (or handed back by the stack) is postponed here: ONE copy of this code per trip
            if (active && ts.node < 0 && nq < MQ) {
                my_q[32 * nq] = ~ts.node; ++nq;
                ts.node = stack[ts.sp]; --ts.sp;
            }
            const bool wn = active && trav_wants_node(ts);
            const unsigned m_node = __ballot_sync(GPT_FULL_MASK, wn);
            const unsigned m_hold = __ballot_sync(GPT_FULL_MASK, active && nq > 0);
            if (m_node == 0u || __popc(m_hold) >= hold_k) break;
            if (wn) {
                ts.node = WIDE ? node4_next(sc.wnodes, ts, stack) : node_next(sc.nodes, ts, stack);      // < 0: a leaf, postponed next trip or stalled on
                if (ts.node == GPT_SENTINEL && nq == 0) {
                    hit[pid] = make_int2(__float_as_int(ts.t), ts.tri);
                    active = false;
                }
            }
        }
        if (__ballot_sync(GPT_FULL_MASK, active) == 0u) break;      // nothing armed and nothing left to fetch
        // ------------------------------------------------------------ leaf phase
        if (active && ts.node < 0) {                                 // the leaf the lane stalled on joins the phase
            my_q[32 * nq] = ~ts.node; ++nq;
            ts.node = stack[ts.sp]; --ts.sp;
        }
        if (!active) nq = 0;
        if (!POOLED) {
            // every lane walks its own leaves, in the order it met them (cudaUtils.h:400-446)
            const float4 ro = my_ray[0], rd = my_ray[1];
            const f3 o = mk3(ro.x, ro.y, ro.z), d = mk3(rd.x, rd.y, rd.z);
            for (int k = 0; k < nq; ++k) {
                const int x = my_q[32 * k];
                int ti = GPT_LEAF_FIRST(x);
                for (;;) {
                    const float4 a = __ldg(sc.tris + 3 * ti), b = __ldg(sc.tris + 3 * ti + 1), c = __ldg(sc.tris + 3 * ti + 2);
                    const float t = tri_test(a, b, c, o, d, CULL);
                    if (t > 0.0f && t < ts.t) { ts.t = t; ts.tri = ti; }
                    if (__float_as_int(c.w) == 0) break;
                    ++ti;
                }
            }
        } else {
            int mine = 0;
            for (int k = 0; k < nq; ++k) mine += pq_leaf_count(sc.tris, my_q[32 * k]);
            int incl = mine;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) { const int v = __shfl_up_sync(GPT_FULL_MASK, incl, off); if ((int)lane >= off) incl += v; }
            const int total = __shfl_sync(GPT_FULL_MASK, incl, 31);
            const int excl = incl - mine;
            if (active) my_ray[0].w = ts.t;              // hitT the pairs are filtered against
            unsigned long long* best = sm.best[warp];
            best[lane] = ~0ull;
            for (int base = 0; base < total; base += WF_PQ_CAP) {
                __syncwarp();
                // lay this chunk's pairs out: each owner writes its own, in the order it met them
                {
                    int g = excl - base;                 // position of the lane's next pair inside the chunk
                    for (int k = 0; k < nq; ++k) {
                        const int x = my_q[32 * k];
                        const int f = GPT_LEAF_FIRST(x), c = pq_leaf_count(sc.tris, x);
#pragma unroll 1
                        for (int i = 0; i < c; ++i, ++g) if ((unsigned)g < (unsigned)WF_PQ_CAP) items[g] = (lane << 27) | (unsigned)(f + i);
                    }
                }
                __syncwarp();
                const int n_chunk = total - base < WF_PQ_CAP ? total - base : WF_PQ_CAP;
#pragma unroll 1
                for (int j = (int)lane; j < n_chunk; j += 32) {
                    const unsigned it = items[j];
                    const unsigned own = it >> 27;
                    const int ti = (int)(it & 0x07ffffffu);
                    const float4 ro = sm.ray[warp][own][0], rd = sm.ray[warp][own][1];
                    const float4 a = __ldg(sc.tris + 3 * ti), b = __ldg(sc.tris + 3 * ti + 1), c = __ldg(sc.tris + 3 * ti + 2);
                    const float t = tri_test(a, b, c, mk3(ro.x, ro.y, ro.z), mk3(rd.x, rd.y, rd.z), CULL);
                    // positive floats order like their bit patterns; ties go to the pair met first
                    if (t > 0.0f && t < ro.w) atomicMin(best + own, ((unsigned long long)__float_as_uint(t) << 32) | (unsigned long long)(unsigned)j);
                }
                __syncwarp();
                const unsigned long long k = best[lane];
                if (k != ~0ull) {                        // strictly nearer than everything tested before: cudaUtils.h:428
                    ts.t = __uint_as_float((unsigned)(k >> 32));
                    ts.tri = (int)(items[(unsigned)(k & 0xffffffffull)] & 0x07ffffffu);
                    my_ray[0].w = ts.t;
                    best[lane] = ~0ull;
                }
            }
        }
        nq = 0;
        if (active && ts.node == GPT_SENTINEL) {
            hit[pid] = make_int2(__float_as_int(ts.t), ts.tri);
            active = false;
        }
    }
}

// One thread per path alive at bounce `bounce`: the rest of the bounce-loop body (tracer.cu:98-296), the path's radiance when
// it ends (tracer.cu:305 / :141, left in accu[pid]), and compaction of the survivors.
// FIRST: bounce 0 -- every path of the wave is alive, the ray is the camera ray of the path id (recomputed, not loaded), the
// mask is 1 (tracer.cu:68) and the accumulated colour starts at 0 (:69), so the slot is WRITTEN here, never read.
// FAST: GPT_FLAG_FAST shading arithmetic (gpt_math.cuh) -- opt-in, statistically equal, not bit-equal.
template <bool FIRST, bool FAST>
__global__ void __launch_bounds__(256, SHADE_MIN_BLOCKS)
k_wf_shade(const DevScene sc, const PrimaryGen gen, const uint32_t bounce,
           const uint32_t* __restrict__ queue_in, unsigned int* __restrict__ counters,
           uint32_t* __restrict__ queue_out, float4* __restrict__ ray_o, float4* __restrict__ ray_d,
           const int2* __restrict__ hit, float* __restrict__ mask_z, float4* __restrict__ accu,
           unsigned long long* __restrict__ stats)
{
    __shared__ SmemPrims sm;
    const float* s_spheres; const float* s_boxes;
    stage_prims(sc, sm, s_spheres, s_boxes);
    const RenderParams& p = gen.p;
    const unsigned n = FIRST ? gen.n_paths : counters[bounce];
    unsigned int* n_out = counters + bounce + 1;
    const unsigned lane = threadIdx.x & 31u;
    const bool has_mesh = sc.n_nodes > 0;
    const bool last_bounce = bounce + 1u >= p.depth;
    const uint32_t n_wave = (uint32_t)gen.n_wave, pix0 = (uint32_t)gen.pix0;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        atomicAdd(&stats[0], (unsigned long long)n);
        if (FIRST) atomicAdd(&stats[1], (unsigned long long)n);
        if ((p.flags & RP_COUNT_BOUNCES) && bounce < 16) atomicAdd(&stats[2 + bounce], (unsigned long long)n);
    }
    // whole warps iterate together so the compaction ballot below is warp-uniform
    const unsigned n_round = (n + 31u) & ~31u;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += gridDim.x * blockDim.x) {
        bool survive = false;
        uint32_t pid = 0;
        if (i < n) {
            PathState ps;
            if (FIRST) {
                pid = i;
                primary_ray(gen, pid, ps);
            } else {
                pid = queue_in[i];
                const float4 ro = ray_o[pid], rd = ray_d[pid];
                ps.o = mk3(ro.x, ro.y, ro.z); ps.d = mk3(rd.x, rd.y, rd.z);
                ps.mask = mk3(ro.w, rd.w, mask_z[pid]);
            }
            ps.accu = mk3(0.0f, 0.0f, 0.0f);      // accu stays in memory (shade_bounce, accu_slot)
            float tri_t = GPT_F32_MAX; int tri = -1;
            if (has_mesh) { const int2 h = hit[pid]; tri_t = __int_as_float(h.x); tri = h.y; }
            const SceneHit h = finish_scene_hit<FAST>(sc, s_spheres, s_boxes, tri_t, tri, ps.o, ps.d);
            const uint32_t s_local = pid / n_wave;
            // accu[pid] is touched inside only on a miss or an emitting hit (always written at bounce 0); when the path is over its
            // radiance (tracer.cu:305, or bkColor on a miss :141) stays there until the wave's samples are summed per pixel
            const bool cont = shade_bounce<FIRST, FAST>(sc, p, s_spheres, s_boxes, h, pix0 + (pid - s_local * n_wave), gen.sample0 + s_local, bounce, ps, accu + pid);
            survive = cont && !last_bounce;
            if (survive) {
                ray_o[pid] = make_float4(ps.o.x, ps.o.y, ps.o.z, ps.mask.x);
                ray_d[pid] = make_float4(ps.d.x, ps.d.y, ps.d.z, ps.mask.y);
                mask_z[pid] = ps.mask.z;
            }
        }
        // ---- compaction: one atomic per warp ----
        const unsigned m = __ballot_sync(GPT_FULL_MASK, survive);
        if (m) {
            unsigned base = 0;
            const int leader = __ffs(m) - 1;
            if ((int)lane == leader) base = atomicAdd(n_out, (unsigned)__popc(m));
            base = __shfl_sync(GPT_FULL_MASK, base, leader);
            if (survive) queue_out[base + (unsigned)__popc(m & ((1u << lane) - 1u))] = pid;
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
k_wf_sort_keys(const DevScene sc, const uint32_t* __restrict__ queue, const unsigned int* __restrict__ n_ptr, const unsigned n_slots,
               const float4* __restrict__ ray_o, const float4* __restrict__ ray_d, uint32_t* __restrict__ keys, const int mode)
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
        if (j < n) {
            const uint32_t pid = queue[j];
            const float4 o = ray_o[pid], d = ray_d[pid];
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
    // extend schedule (GPT_WF_EXTEND): ww = the plain while-while loop (k_wf_extend_ww); p3 / p1 = pooled leaf phases with three /
    // one postponed leaves per lane; r1 = in-loop re-arming with per-lane leaf phases (k_wf_extend_pq)
    { const char* x = getenv("GPT_WF_EXTEND"); cfg->pooled = !x ? WF_DEFAULT_SCHEDULE : !strcmp(x, "p3") ? 3 : !strcmp(x, "p1") ? 1 : !strcmp(x, "r1") ? 2 : 0; }
    cfg->pq_grid = sm_count * per_sm(cfg->pooled == 3 ? (const void*)k_wf_extend_pq<true, false, false, 3, true> : cfg->pooled == 1 ? (const void*)k_wf_extend_pq<true, false, false, 1, true> : (const void*)k_wf_extend_pq<true, false, false, 1, false>, WF_THREADS, 1);
    cfg->pq_wide_grid = sm_count * per_sm(cfg->pooled == 3 ? (const void*)k_wf_extend_pq<true, true, false, 3, true> : cfg->pooled == 1 ? (const void*)k_wf_extend_pq<true, true, false, 1, true> : (const void*)k_wf_extend_pq<true, true, false, 1, false>, WF_THREADS, 1);
    { const char* h = getenv("GPT_WF_HOLD_K"); cfg->hold_k_env = (h && atoi(h) > 0) ? wf_knob("GPT_WF_HOLD_K", 20) : 0; }      // 0 / unset: by scene size
    cfg->pq_refill_min = wf_knob("GPT_WF_PQ_REFILL_MIN", 4);
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
                          const WaveBuffers& wb, int node_k, const PrimaryGen& gen, cudaStream_t st)
{
    if (cfg.pooled) {
        // leaves held in the warp that start a leaf phase: pooled phases want them full, per-lane phases the reference's few stragglers
        const int hold_k = cfg.hold_k_env ? cfg.hold_k_env : (cfg.pooled == 3 ? 20 : 16);
        const int grid = sc.wnodes ? cfg.pq_wide_grid : cfg.pq_grid;
#define PQ_LAUNCH(C, Wd, M, P) k_wf_extend_pq<C, Wd, PRIMARY, M, P><<<grid, WF_THREADS, 0, st>>>(sc, q, n_ptr, fetch, wb.ray_o, wb.ray_d, wb.hit, cfg.pq_refill_min, hold_k, gen)
#define PQ_PICK(M, P) { if (sc.wnodes) { if (p.cull) PQ_LAUNCH(true, true, M, P); else PQ_LAUNCH(false, true, M, P); } \
                        else { if (p.cull) PQ_LAUNCH(true, false, M, P); else PQ_LAUNCH(false, false, M, P); } }
        if (cfg.pooled == 3) PQ_PICK(3, true)
        else if (cfg.pooled == 1) PQ_PICK(1, true)
        else PQ_PICK(1, false)
#undef PQ_PICK
#undef PQ_LAUNCH
        return;
    }
    if (sc.wnodes) {
        if (p.cull) k_wf_extend_ww<true, true, PRIMARY><<<cfg.wide_grid, WF_THREADS, 0, st>>>(sc, q, n_ptr, fetch, wb.ray_o, wb.ray_d, wb.hit, cfg.refill_min, node_k, gen);
        else k_wf_extend_ww<false, true, PRIMARY><<<cfg.wide_grid, WF_THREADS, 0, st>>>(sc, q, n_ptr, fetch, wb.ray_o, wb.ray_d, wb.hit, cfg.refill_min, node_k, gen);
    } else {
        if (p.cull) k_wf_extend_ww<true, false, PRIMARY><<<cfg.ext_grid, WF_THREADS, 0, st>>>(sc, q, n_ptr, fetch, wb.ray_o, wb.ray_d, wb.hit, cfg.refill_min, node_k, gen);
        else k_wf_extend_ww<false, false, PRIMARY><<<cfg.ext_grid, WF_THREADS, 0, st>>>(sc, q, n_ptr, fetch, wb.ray_o, wb.ray_d, wb.hit, cfg.refill_min, node_k, gen);
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
    const int node_k = cfg.node_k_env ? cfg.node_k_env : (sc.n_tris < (64 << 10) ? WF_NODE_K : (sc.n_tris < (2 << 20) ? 8 : 16));      // 8: between the reference-built dragon's best (6) and the device-built trees' (10)
    const int n_pix = p.width * p.height;
    const int n_counters = (int)gptk_wave_counter_count(p.depth);
    // A wave = a run of consecutive pixels x a run of consecutive samples, as many paths as the scratch arrays
    // hold (68 B/path).  Big waves matter: every launch has a fixed cost (launch gap + the tail where the
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
                const uint32_t* q_in = b == 0 ? nullptr : wb.queue[b & 1u];
                const uint32_t* q_ext = q_in;       // order in which the extend kernel fetches the rays
                if (has_mesh && b > 0 && wb.sort_vals) {
                    if (kt) kt->tick(KT_OTHER, st);
                    k_wf_sort_keys<<<grid, 256, 0, st>>>(sc, q_in, wb.counters + b, (unsigned)n_paths, wb.ray_o, wb.ray_d, wb.sort_keys[0], cfg.sort_mode);
                    size_t tb = wb.sort_tmp_bytes;
                    if ((e = cub::DeviceRadixSort::SortPairs(wb.sort_tmp, tb, (const uint32_t*)wb.sort_keys[0], wb.sort_keys[1], q_in, wb.sort_vals,
                                                             (long long)n_paths, 0, wf_sort_bits(cfg.sort_mode), st)) != cudaSuccess) return e;
                    *launches += 4;
                    q_ext = wb.sort_vals;
                }
                if (has_mesh) {
                    if (kt) kt->tick(KT_EXTEND, st);
                    if (b == 0) launch_extend<true>(sc, p, cfg, nullptr, wb.counters, wb.counters + p.depth + 1, wb, node_k, gen, st);
                    else launch_extend<false>(sc, p, cfg, q_ext, wb.counters + b, wb.counters + p.depth + 1 + b, wb, node_k, gen, st);
                    *launches += 1;
                }
                if (kt) kt->tick(KT_SHADE, st);
                if (p.flags & RP_FAST) {
                    if (b == 0) k_wf_shade<true, true><<<sgrid, 256, 0, st>>>(sc, gen, b, nullptr, wb.counters, wb.queue[1], wb.ray_o, wb.ray_d, wb.hit, wb.mask_z, wb.accu, stats);
                    else k_wf_shade<false, true><<<sgrid, 256, 0, st>>>(sc, gen, b, q_in, wb.counters, wb.queue[(b + 1u) & 1u], wb.ray_o, wb.ray_d, wb.hit, wb.mask_z, wb.accu, stats);
                } else {
                    if (b == 0) k_wf_shade<true, false><<<sgrid, 256, 0, st>>>(sc, gen, b, nullptr, wb.counters, wb.queue[1], wb.ray_o, wb.ray_d, wb.hit, wb.mask_z, wb.accu, stats);
                    else k_wf_shade<false, false><<<sgrid, 256, 0, st>>>(sc, gen, b, q_in, wb.counters, wb.queue[(b + 1u) & 1u], wb.ray_o, wb.ray_d, wb.hit, wb.mask_z, wb.accu, stats);
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
