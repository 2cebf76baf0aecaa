// gpt_traverse.cuh -- nearest-hit BVH2 traversal + ray/triangle/sphere/box intersection.
//
// Replaces intersectBVHandTriangles (cudaUtils.h:256-460), intersectRayTriangleEdge (:135-172),
// spanBegin/EndKepler (:239-250), intersectAllSpeheres (:221-236), Sphere::intersect
// (CommomStructs.hpp:23-31) and the Box spec (cudaUtils.h:31-77).
//
// Same algorithmic contract as the reference (depth-first, near child first by entry distance,
// strict t < hitT so the first-visited triangle wins ties), different mechanics:
//   * one 64-B node = 4 x LDG.128 through the read-only path issued together (the reference
//     fetches the child pointers with a dependent second load),
//   * triangles are pre-packed (v0, e1, e2) with the original index / last-in-leaf flag in the w
//     lanes, so a leaf costs 48 B per triangle and no terminator or index-array read,
//   * FMNMX min/max instead of the Kepler video instructions,
//   * leaf postponing voted with __ballot_sync over the currently converged lanes.
#pragma once
#include "gpt_scene.cuh"

#ifndef SPHERE_UNROLL
#define SPHERE_UNROLL 3
#endif
struct TriHit { float t; int tri; };   // tri = packed triangle index, -1 = none

// Moller-Trumbore on a packed triangle; formulas and test order of cudaUtils.h:143-169.
// Returns t, or a negative value for a miss.
__device__ __forceinline__ float tri_test(const float4 a, const float4 b, const float4 c, const f3 o, const f3 d, const bool cull)
{
    const f3 v0 = mk3(a.x, a.y, a.z), e1 = mk3(b.x, b.y, b.z), e2 = mk3(c.x, c.y, c.z);
    const f3 pvec = cross3(d, e2);
    const float det = dot3(e1, pvec);
    if (det < -0.00001f) { if (cull) return -1.0f; }
    else if (det < 0.00001f) return -1.0f;
    const float invdet = 1.0f / det;
    const f3 tvec = sub3(o, v0);
    const float u = dot3(tvec, pvec) * invdet;
    if (u < 0.0f || u > 1.0f) return -1.0f;
    const f3 qvec = cross3(tvec, e1);
    const float v = dot3(d, qvec) * invdet;
    if (v < 0.0f || (u + v) > 1.0f) return -1.0f;
    const float t = dot3(e2, qvec) * invdet;
    return (t > 0.0f && t < GPT_F32_MAX) ? t : -1.0f;
}

__device__ __forceinline__ float guard_dir(float d)
{
    const float ooeps = 8.2718061255302767e-25f;   // 2^-80, cudaUtils.h:283
    return fabsf(d) > ooeps ? d : copysignf(ooeps, d);
}

// Per-lane traversal state.  Two ways to drive it:
//   * traverse_bvh(): the classic while-while loop (inner nodes until the converged lanes all hold a leaf,
//     then the leaves) run to completion -- used by the test hooks;
//   * node_step()/tri_step(): ONE inner-node visit or ONE triangle test per call -- used by the render
//     megakernel, whose warps vote each trip on which kind of step to run (see gpt_kernels.cu).
// Both visit nodes and triangles of a ray in the same depth-first, near-child-first order as the reference
// (cudaUtils.h:298-448), so (hitT, hit triangle) -- including ties -- are identical.
struct TravState {
    float idx, idy, idz;      // 1/dir (2^-80 guarded, cudaUtils.h:283-286)
    float oodx, oody, oodz;   // origin/dir
    float t;                  // closest hit so far (hitT)
    int   tri;                // packed triangle index of that hit, -1 = none
    int   node;               // float4 index of the current inner node; <0 a second leaf; SENTINEL = stack empty
    int   leaf;               // postponed leaf (<0) or none (>=0)
    int   tri_cur;            // next triangle of the postponed leaf to test
    int   sp;
};

__device__ __forceinline__ void trav_init(TravState& s, int* stack, const f3 o, const f3 d, const bool has_mesh)
{
    s.idx = 1.0f / guard_dir(d.x); s.idy = 1.0f / guard_dir(d.y); s.idz = 1.0f / guard_dir(d.z);
    s.oodx = o.x * s.idx; s.oody = o.y * s.idy; s.oodz = o.z * s.idz;
    s.t = GPT_F32_MAX; s.tri = -1;
    s.sp = 0; stack[0] = GPT_SENTINEL;
    s.node = has_mesh ? 0 : GPT_SENTINEL;    // root = float4 index 0
    s.leaf = 0; s.tri_cur = 0;
}

__device__ __forceinline__ bool trav_wants_node(const TravState& s) { return (unsigned)s.node < (unsigned)GPT_SENTINEL; }
__device__ __forceinline__ bool trav_wants_tri(const TravState& s) { return s.leaf < 0; }
__device__ __forceinline__ bool trav_finished(const TravState& s) { return s.node == GPT_SENTINEL && s.leaf >= 0; }

// One inner-node visit, cudaUtils.h:305-360: the child to go to next (inner node >= 0, leaf < 0, or what the stack gives back),
// the far child pushed.  Requires trav_wants_node(s).
__device__ __forceinline__ int node_next(const float4* __restrict__ nodes, TravState& s, int* stack)
{
    const int node = s.node;
    float4 n0, n1, nz, chf;
    ldg_node64(nodes + node, n0, n1, nz, chf);
    const int4 ch = make_int4(__float_as_int(chf.x), __float_as_int(chf.y), 0, 0);
    const float c0lox = fmaf(n0.x, s.idx, -s.oodx), c0hix = fmaf(n0.y, s.idx, -s.oodx);
    const float c0loy = fmaf(n0.z, s.idy, -s.oody), c0hiy = fmaf(n0.w, s.idy, -s.oody);
    const float c1lox = fmaf(n1.x, s.idx, -s.oodx), c1hix = fmaf(n1.y, s.idx, -s.oodx);
    const float c1loy = fmaf(n1.z, s.idy, -s.oody), c1hiy = fmaf(n1.w, s.idy, -s.oody);
    const float c0loz = fmaf(nz.x, s.idz, -s.oodz), c0hiz = fmaf(nz.y, s.idz, -s.oodz);
    const float c1loz = fmaf(nz.z, s.idz, -s.oodz), c1hiz = fmaf(nz.w, s.idz, -s.oodz);
    // tmin = 0 (tracer.cu:94): max(...,0) ; tmax = current hitT
    const float c0min = fmaxf(fmaxf(fminf(c0lox, c0hix), fminf(c0loy, c0hiy)), fmaxf(fminf(c0loz, c0hiz), 0.0f));
    const float c0max = fminf(fminf(fmaxf(c0lox, c0hix), fmaxf(c0loy, c0hiy)), fminf(fmaxf(c0loz, c0hiz), s.t));
    const float c1min = fmaxf(fmaxf(fminf(c1lox, c1hix), fminf(c1loy, c1hiy)), fmaxf(fminf(c1loz, c1hiz), 0.0f));
    const float c1max = fminf(fminf(fmaxf(c1lox, c1hix), fmaxf(c1loy, c1hiy)), fminf(fmaxf(c1loz, c1hiz), s.t));
    // cudaUtils.h:333-334: (min<=max) && (min>=tmin) && (min<=F32_MAX)
    const bool t0 = (c0min <= c0max) && (c0min <= GPT_F32_MAX);
    const bool t1 = (c1min <= c1max) && (c1min <= GPT_F32_MAX);
    int next;
    if (!t0 && !t1) {
        next = stack[s.sp]; --s.sp;
    } else {
        next = t0 ? ch.x : ch.y;
        if (t0 && t1) {
            int far_child = ch.y;
            if (c1min < c0min) { far_child = next; next = ch.y; }
            ++s.sp; stack[s.sp] = far_child;
        }
    }
    return next;
}

// One inner-node visit with the reference's leaf handling, cudaUtils.h:305-369.
__device__ __forceinline__ void node_step(const float4* __restrict__ nodes, TravState& s, int* stack)
{
    int next = node_next(nodes, s, stack);
    if (next < 0 && s.leaf >= 0) {           // first leaf: postpone it and keep descending (:362-369)
        s.leaf = next; s.tri_cur = ~next;
        next = stack[s.sp]; --s.sp;
    }
    s.node = next;
}

// One visit of a 4-wide quantised node (gpt_wide.cu): same state machine as node_step -- nearest hit child next, the others
// pushed farthest first, a first leaf postponed -- over four children per 64-B fetch instead of two.
// 0x4B000000 = bits of 2^23.  A __constant__ the compiler cannot fold: as a literal it would be re-materialised as a second
// immediate next to every PRMT selector and cost a register move per plane (measured: 24 fewer instructions per visit, +1 %;
// masking the bytes in with LOP3 at two scales instead was 2 % slower).
static __constant__ unsigned gpt_w4_magic = 0x4B000000u;
#define W4_CSWAP(ka, ca, kb, cb) { const bool sw_ = kb < ka; const float tk_ = sw_ ? ka : kb; const int tc_ = sw_ ? ca : cb; ka = sw_ ? kb : ka; ca = sw_ ? cb : ca; kb = tk_; cb = tc_; }
__device__ __forceinline__ int node4_next(const float4* __restrict__ wn, TravState& s, int* stack)
{
    const int node = s.node;
    float4 h, qaf, qbf, chf;
    ldg_node64(wn + node, h, qaf, qbf, chf);
    const uint4 qa = make_uint4(__float_as_uint(qaf.x), __float_as_uint(qaf.y), __float_as_uint(qaf.z), __float_as_uint(qaf.w));
    const uint2 qb = make_uint2(__float_as_uint(qbf.x), __float_as_uint(qbf.y));
    const int4 ch = make_int4(__float_as_int(chf.x), __float_as_int(chf.y), __float_as_int(chf.z), __float_as_int(chf.w));
    const unsigned eb = __float_as_uint(h.w);
    // plane = origin + q * 2^e  =>  t = (plane - o) / d = q * s + b  with s = 2^e / d, b = (origin - o) / d.  The byte q is
    // dropped into the mantissa of 2^23 (one PRMT, selector immediate, 2^23 kept in a register) and the 2^23 * s it
    // brings along is taken out of b once per axis; gpt_wide.cu leaves a cell of slack for the rounding of that.
    const unsigned magic = gpt_w4_magic;     // bits of 2^23, read from constant memory so that it stays in ONE register
    const float sx = __uint_as_float((eb & 0xffu) << 23) * s.idx;
    const float sy = __uint_as_float((eb & 0xff00u) << 15) * s.idy;
    const float sz = __uint_as_float((eb & 0xff0000u) << 7) * s.idz;
    const float bx = fmaf(-8388608.0f, sx, fmaf(h.x, s.idx, -s.oodx));
    const float by = fmaf(-8388608.0f, sy, fmaf(h.y, s.idy, -s.oody));
    const float bz = fmaf(-8388608.0f, sz, fmaf(h.z, s.idz, -s.oodz));
    // entry planes are the lo planes where the direction is positive, the hi planes otherwise
    const bool px = s.idx >= 0.0f, py = s.idy >= 0.0f, pz = s.idz >= 0.0f;
    const unsigned nx = px ? qa.x : qa.y, fx = px ? qa.y : qa.x;
    const unsigned ny = py ? qa.z : qa.w, fy = py ? qa.w : qa.z;
    const unsigned nz = pz ? qb.x : qb.y, fz = pz ? qb.y : qb.x;
    const float inf = __int_as_float(0x7f800000);
    float k0, k1, k2, k3;
    int c0 = ch.x, c1 = ch.y, c2 = ch.z, c3 = ch.w;
    // an unused slot has an inverted box (entry plane beyond exit plane on every axis): tn > tf without a test of its own
#define W4_Q(w, i) __uint_as_float(__byte_perm(w, magic, 0x7650u | (i)))
#define W4_CHILD(i, kk) { \
        const float tn = fmaxf(fmaxf(fmaf(W4_Q(nx, i), sx, bx), fmaf(W4_Q(ny, i), sy, by)), fmaxf(fmaf(W4_Q(nz, i), sz, bz), 0.0f)); \
        const float tf = fminf(fminf(fmaf(W4_Q(fx, i), sx, bx), fmaf(W4_Q(fy, i), sy, by)), fminf(fmaf(W4_Q(fz, i), sz, bz), s.t)); \
        kk = tn <= tf ? tn : inf; }
    W4_CHILD(0u, k0) W4_CHILD(1u, k1) W4_CHILD(2u, k2) W4_CHILD(3u, k3)
#undef W4_CHILD
#undef W4_Q
    W4_CSWAP(k0, c0, k1, c1) W4_CSWAP(k2, c2, k3, c3) W4_CSWAP(k0, c0, k2, c2) W4_CSWAP(k1, c1, k3, c3) W4_CSWAP(k1, c1, k2, c2)
    int next;
    if (!(k0 < inf)) {
        next = stack[s.sp]; --s.sp;
    } else {
        next = c0;
        if (k3 < inf) { ++s.sp; stack[s.sp] = c3; }
        if (k2 < inf) { ++s.sp; stack[s.sp] = c2; }
        if (k1 < inf) { ++s.sp; stack[s.sp] = c1; }
    }
    return next;
}
__device__ __forceinline__ void node4_step(const float4* __restrict__ wn, TravState& s, int* stack)
{
    int next = node4_next(wn, s, stack);
    if (next < 0 && s.leaf >= 0) {           // first leaf: postpone it and keep descending
        s.leaf = next; s.tri_cur = ~next;
        next = stack[s.sp]; --s.sp;
    }
    s.node = next;
}

// One triangle of the postponed leaf: cudaUtils.h:405-446.  Requires trav_wants_tri(s).
__device__ __forceinline__ void tri_step(const float4* __restrict__ tris, TravState& s, int* stack,
                                         const f3 o, const f3 d, const bool cull)
{
    const int ti = s.tri_cur;
    float4 b, c;
    tri_rows_bc(tris, ti, b, c);
    const float4 a = tri_row_a(tris, ti);
    const float t = tri_test(a, b, c, o, d, cull);
    if (t > 0.0f && t < s.t) { s.t = t; s.tri = ti; }        // :428, strict: first visited wins ties
    if (__float_as_int(c.w) == 0) {                          // last triangle of the leaf (no triangle follows)
        s.leaf = s.node;                                     // :440 another leaf was reached meanwhile?
        if (s.node < 0) { s.tri_cur = ~s.node; s.node = stack[s.sp]; --s.sp; }
    } else {
        s.tri_cur = ti + 1;
    }
}

// A whole postponed leaf with back-face culling on, in two passes (same tests, same order of acceptance as tri_step):
//   A: only the determinant test of cudaUtils.h:145-153 for every triangle of the leaf -- with culling it rejects about half
//      of them after 2 of the 3 loads and ~12 of the ~90 instructions -- survivors recorded in a bit mask;
//   B: the full Moller-Trumbore test (tri_test, which recomputes the same determinant) for the survivors, in leaf order.
// In tri_step's one-triangle-per-trip loop a lane whose triangle fails the determinant test still waits for the lanes that
// go on; here the expensive pass runs max(survivors) trips per warp instead of max(leaf size).
__device__ __forceinline__ void leaf_step_culled(const float4* __restrict__ tris, TravState& s, int* stack, const f3 o, const f3 d)
{
    int first = s.tri_cur;
    bool more = true;
    while (more) {
        unsigned m = 0u; int k = 0;
        do {
            float4 b, c;
            tri_rows_bc(tris, first + k, b, c);
            const float det = dot3(mk3(b.x, b.y, b.z), cross3(d, mk3(c.x, c.y, c.z)));
            if (!(det < 0.00001f)) m |= 1u << k;             // culling: det < -eps and -eps <= det < eps both miss
            more = __float_as_int(c.w) != 0;                 // triangles that follow in this leaf
            ++k;
        } while (more && k < 32);
        while (m) {
            const int ti = first + __ffs(m) - 1;
            m &= m - 1u;
            float4 b, c;
            tri_rows_bc(tris, ti, b, c);
            const float t = tri_test(tri_row_a(tris, ti), b, c, o, d, true);
            if (t > 0.0f && t < s.t) { s.t = t; s.tri = ti; }   // :428, strict: first visited wins ties
        }
        first += k;
    }
    s.leaf = s.node;                                         // :440 another leaf was reached meanwhile?
    if (s.node < 0) { s.tri_cur = ~s.node; s.node = stack[s.sp]; --s.sp; }
}

__device__ __forceinline__ TriHit traverse_bvh(const float4* __restrict__ nodes, const float4* __restrict__ tris,
                                               const f3 o, const f3 d, const bool cull)
{
    int stack[GPT_STACK_SIZE];
    TravState s;
    trav_init(s, stack, o, d, true);
    while (!trav_finished(s)) {
        // inner nodes until every converged lane holds a leaf (or ran out of nodes)
        while (trav_wants_node(s)) {
            node_step(nodes, s, stack);
            if (!__any_sync(__activemask(), s.leaf >= 0)) break;
        }
        while (trav_wants_tri(s)) tri_step(tris, s, stack, o, d, cull);
    }
    TriHit hit; hit.t = s.t; hit.tri = s.tri;
    return hit;
}

// same loop over the 4-wide tree
__device__ __forceinline__ TriHit traverse_bvh4(const float4* __restrict__ wnodes, const float4* __restrict__ tris,
                                                const f3 o, const f3 d, const bool cull)
{
    int stack[GPT_W4_STACK];
    TravState s;
    trav_init(s, stack, o, d, true);
    while (!trav_finished(s)) {
        while (trav_wants_node(s)) {
            node4_step(wnodes, s, stack);
            if (!__any_sync(__activemask(), s.leaf >= 0)) break;
        }
        while (trav_wants_tri(s)) tri_step(tris, s, stack, o, d, cull);
    }
    TriHit hit; hit.t = s.t; hit.tri = s.tri;
    return hit;
}

// Sphere::intersect, CommomStructs.hpp:23-31; s = 11 floats (posRad, emi, col, mat)
template <bool FAST = false>
__device__ __forceinline__ float sphere_test(const float* s, const f3 o, const f3 d)
{
    const f3 op = sub3(mk3(s[0], s[1], s[2]), o);
    const float b = dot3(op, d);
    const float disc = fmaf(s[3], s[3], fmaf(b, b, -dot3(op, op)));      // contraction as in the reference's SASS
    // branch-free: the warp pays for the square root whenever any lane needs it, so every lane may as well
    // take it; the selected values are the same as in the early-out form
    const float sq = sqrt_t<FAST>(fmaxf(disc, 0.0f));
    const float t0 = b - sq, t1 = b + sq;
    const float t = t0 > 0.01f ? t0 : (t1 > 0.01f ? t1 : 0.0f);
    return disc < 0.0f ? 0.0f : t;
}

// Box spec, cudaUtils.h:43-58; bx = 13 floats (min, max, emi, col, mat)
__device__ __forceinline__ float box_test(const float* bx, const f3 o, const f3 d)
{
    const float t0x = (bx[0] - o.x) / d.x, t1x = (bx[3] - o.x) / d.x;
    const float t0y = (bx[1] - o.y) / d.y, t1y = (bx[4] - o.y) / d.y;
    const float t0z = (bx[2] - o.z) / d.z, t1z = (bx[5] - o.z) / d.z;
    const float minmax = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fmaxf(t0z, t1z));
    const float maxmin = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fminf(t0z, t1z));
    if (minmax >= maxmin) return maxmin > 0.001f ? maxmin : 0.0f;
    return 0.0f;
}

// Full nearest hit of one bounce: BVH, then the sphere sweep (cudaUtils.h:221-236: a sphere
// replaces the triangle hit only if strictly closer and > 0.01), then boxes.
struct SceneHit { float t; int geom; int prim; };   // prim = packed tri / sphere / box index

// sphere sweep (cudaUtils.h:221-236) + boxes on top of a finished BVH result
template <bool FAST = false>
__device__ __forceinline__ SceneHit finish_scene_hit(const DevScene& sc, const float* s_spheres, const float* s_boxes,
                                                     const float tri_t, const int tri, const f3 o, const f3 d)
{
    SceneHit h; h.t = tri_t; h.geom = GEO_NONE; h.prim = -1;
    if (tri >= 0) { h.geom = GEO_TRI; h.prim = tri; }
    // three tests in flight: the dependent chain of one test (dot, fma, MUFU.RSQ, fix-up) no longer stalls the warp, and the
    // shade kernel needs fewer registers that way (52 instead of 58; 46 under its 5-CTA bound, no spills)
    constexpr int sphere_unroll = SPHERE_UNROLL;
#pragma unroll sphere_unroll
    for (int i = 0; i < sc.n_spheres; ++i) {
        const float t = sphere_test<FAST>(s_spheres + 11 * i, o, d);
        if (t != 0.0f && t < h.t && t > 0.01f) { h.t = t; h.geom = GEO_SPHERE; h.prim = i; }
    }
    for (int i = 0; i < sc.n_boxes; ++i) {
        const float t = box_test(s_boxes + 13 * i, o, d);
        if (t != 0.0f && t < h.t) { h.t = t; h.geom = GEO_BOX; h.prim = i; }
    }
    return h;
}

// the same sweep with the spheres' centre + radius read as one float4 each (an array in shared memory: LDS.128); same arithmetic
template <bool FAST = false>
__device__ __forceinline__ SceneHit finish_scene_hit_smem(const DevScene& sc, const float4* posrad, const float* s_boxes,
                                                          const float tri_t, const int tri, const f3 o, const f3 d)
{
    SceneHit h; h.t = tri_t; h.geom = GEO_NONE; h.prim = -1;
    if (tri >= 0) { h.geom = GEO_TRI; h.prim = tri; }
    constexpr int sphere_unroll = SPHERE_UNROLL;
#pragma unroll sphere_unroll
    for (int i = 0; i < sc.n_spheres; ++i) {
        const float4 pr = posrad[i];
        const float s4[4] = {pr.x, pr.y, pr.z, pr.w};
        const float t = sphere_test<FAST>(s4, o, d);
        if (t != 0.0f && t < h.t && t > 0.01f) { h.t = t; h.geom = GEO_SPHERE; h.prim = i; }
    }
    for (int i = 0; i < sc.n_boxes; ++i) {
        const float t = box_test(s_boxes + 13 * i, o, d);
        if (t != 0.0f && t < h.t) { h.t = t; h.geom = GEO_BOX; h.prim = i; }
    }
    return h;
}

__device__ __forceinline__ SceneHit intersect_scene(const DevScene& sc, const float* s_spheres, const float* s_boxes,
                                                    const f3 o, const f3 d, const bool cull)
{
    TriHit th; th.t = GPT_F32_MAX; th.tri = -1;
    if (sc.n_nodes > 0) th = traverse_bvh(sc.nodes, sc.tris, o, d, cull);
    return finish_scene_hit(sc, s_spheres, s_boxes, th.t, th.tri, o, d);
}
