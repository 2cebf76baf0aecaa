// gpt_build.cu -- device-side BVH builder.
//
// The reference builds its tree on ONE host thread (SplitBVHBuilder.cpp:105-164: three full sorts per node) and
// flattens it with CudaBVH::createCompact (CudaBVH.cpp:121-270): 2.9 s per 100 k triangles, minutes for the
// 20 M-triangle configuration.  This builder runs on the GPU and emits the SAME array layout createCompact emits
// (node array 4 x vec4 per inner node, raw-vertex triangle array with one 0x80000000 terminator per leaf, index array
// -- CudaBVH.cpp:196-224), so its output is a drop-in for kernelInfo.bvhData / gpt_upload_scene and the CPU oracle
// can traverse the very same arrays.  The tree is NOT the reference's SBVH (no spatial splits): it is
//   1. 63-bit Morton codes of the triangle AABB centres, radix-sorted (CUB),
//   2. a topology over the sorted triangles: PLOC clustering (ploc_radius > 0, the default -- see k_ploc_* below) or the
//      binary radix tree of Karras 2012 over the sorted codes (ties broken by position),
//   3. bottom-up AABB fit with a SAH evaluation (node cost 1, triangle cost 1 -- the reference's Platform() defaults,
//      bvh_util.hpp) that collapses a subtree into one leaf when  n * A(node) <= A(node) + cost(left) + cost(right)
//      and n <= max_leaf,
//   4. a level-synchronous top-down numbering (children always behind their parent, like createCompact),
//   5. emission of the three arrays.
// Nearest-hit distances are independent of the tree (same per-triangle arithmetic); only which of two triangles
// with bit-identical t is reported may differ from a reference-built tree.
#include "gpt_kernels.h"
#include <stdlib.h>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

namespace {

__device__ __forceinline__ int f2ord(float f) { const int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__host__ __device__ __forceinline__ float ord2f(int i) { union { int i; float f; } u; u.i = i >= 0 ? i : i ^ 0x7fffffff; return u.f; }

__global__ void __launch_bounds__(256)
k_prim_bounds(const float* __restrict__ verts, const int n_verts, const int* __restrict__ tri, const int n,
              float4* __restrict__ lo, float4* __restrict__ hi, int* __restrict__ scene, int* __restrict__ bad)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float cx0 = 3.0e38f, cy0 = 3.0e38f, cz0 = 3.0e38f, cx1 = -3.0e38f, cy1 = -3.0e38f, cz1 = -3.0e38f;
    if (i < n) {
        int a = tri[3 * i], b = tri[3 * i + 1], c = tri[3 * i + 2];
        if ((unsigned)a >= (unsigned)n_verts || (unsigned)b >= (unsigned)n_verts || (unsigned)c >= (unsigned)n_verts) { atomicExch(bad, 1); a = b = c = 0; }
        const float ax = verts[3 * a], ay = verts[3 * a + 1], az = verts[3 * a + 2];
        const float bx = verts[3 * b], by = verts[3 * b + 1], bz = verts[3 * b + 2];
        const float cx = verts[3 * c], cy = verts[3 * c + 1], cz = verts[3 * c + 2];
        const float4 l = make_float4(fminf(ax, fminf(bx, cx)), fminf(ay, fminf(by, cy)), fminf(az, fminf(bz, cz)), 0.0f);
        const float4 h = make_float4(fmaxf(ax, fmaxf(bx, cx)), fmaxf(ay, fmaxf(by, cy)), fmaxf(az, fmaxf(bz, cz)), 0.0f);
        // fminf/fmaxf drop NaNs, so test the coordinates themselves
        if (!isfinite(ax + ay + az + bx + by + bz + cx + cy + cz) || !(isfinite(l.x) && isfinite(l.y) && isfinite(l.z) && isfinite(h.x) && isfinite(h.y) && isfinite(h.z))) atomicExch(bad, 2);
        lo[i] = l; hi[i] = h;
        cx0 = cx1 = 0.5f * (l.x + h.x); cy0 = cy1 = 0.5f * (l.y + h.y); cz0 = cz1 = 0.5f * (l.z + h.z);
    }
    for (int o = 16; o > 0; o >>= 1) {
        cx0 = fminf(cx0, __shfl_xor_sync(0xffffffffu, cx0, o)); cy0 = fminf(cy0, __shfl_xor_sync(0xffffffffu, cy0, o));
        cz0 = fminf(cz0, __shfl_xor_sync(0xffffffffu, cz0, o)); cx1 = fmaxf(cx1, __shfl_xor_sync(0xffffffffu, cx1, o));
        cy1 = fmaxf(cy1, __shfl_xor_sync(0xffffffffu, cy1, o)); cz1 = fmaxf(cz1, __shfl_xor_sync(0xffffffffu, cz1, o));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(scene + 0, f2ord(cx0)); atomicMin(scene + 1, f2ord(cy0)); atomicMin(scene + 2, f2ord(cz0));
        atomicMax(scene + 3, f2ord(cx1)); atomicMax(scene + 4, f2ord(cy1)); atomicMax(scene + 5, f2ord(cz1));
    }
}

__device__ __forceinline__ unsigned long long spread21(unsigned int v)
{
    unsigned long long x = v & 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

__global__ void __launch_bounds__(256)
k_morton(const float4* __restrict__ lo, const float4* __restrict__ hi, const int n, const int* __restrict__ scene,
         unsigned long long* __restrict__ keys, int* __restrict__ vals)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float mx = ord2f(scene[0]), my = ord2f(scene[1]), mz = ord2f(scene[2]);
    const float ex = ord2f(scene[3]) - mx, ey = ord2f(scene[4]) - my, ez = ord2f(scene[5]) - mz;
    const float4 l = lo[i], h = hi[i];
    const float qx = ex > 0.0f ? (0.5f * (l.x + h.x) - mx) / ex : 0.0f;
    const float qy = ey > 0.0f ? (0.5f * (l.y + h.y) - my) / ey : 0.0f;
    const float qz = ez > 0.0f ? (0.5f * (l.z + h.z) - mz) / ez : 0.0f;
    const unsigned int ix = (unsigned int)fminf(fmaxf(qx * 2097152.0f, 0.0f), 2097151.0f);
    const unsigned int iy = (unsigned int)fminf(fmaxf(qy * 2097152.0f, 0.0f), 2097151.0f);
    const unsigned int iz = (unsigned int)fminf(fmaxf(qz * 2097152.0f, 0.0f), 2097151.0f);
    keys[i] = (spread21(ix) << 2) | (spread21(iy) << 1) | spread21(iz);
    vals[i] = i;
}

// common-prefix length of sorted keys i and j (Karras 2012, section 4); equal keys fall back to the positions
__device__ __forceinline__ int delta(const unsigned long long* __restrict__ keys, const int n, const int i, const int j)
{
    if (j < 0 || j >= n) return -1;
    const unsigned long long a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz(i ^ j);
    return __clzll((long long)(a ^ b));
}

// internal node i of the radix tree: children (>= 0 internal node, < 0 = ~sorted position of a primitive),
// covered range [first, last] of sorted positions, parent links
__global__ void __launch_bounds__(256)
k_radix_tree(const unsigned long long* __restrict__ keys, const int n, int* __restrict__ child_l, int* __restrict__ child_r,
             int* __restrict__ first, int* __restrict__ last, int* __restrict__ parent_i, int* __restrict__ parent_p)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
    const int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(keys, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + (d < 0 ? d : 0);
    const int lo_ = i < j ? i : j, hi_ = i < j ? j : i;
    const int cl = (lo_ == gamma) ? ~gamma : gamma;
    const int cr = (hi_ == gamma + 1) ? ~(gamma + 1) : gamma + 1;
    child_l[i] = cl; child_r[i] = cr; first[i] = lo_; last[i] = hi_;
    if (cl < 0) parent_p[~cl] = i; else parent_i[cl] = i;
    if (cr < 0) parent_p[~cr] = i; else parent_i[cr] = i;
    if (i == 0) parent_i[0] = -1;
}

__device__ __forceinline__ float half_area(const float4 l, const float4 h)
{
    const float dx = h.x - l.x, dy = h.y - l.y, dz = h.z - l.z;
    return dx * dy + dy * dz + dz * dx;
}

// One thread per sorted primitive climbs towards the root; the second thread to reach a node owns it (its
// children are complete), fits its box and evaluates the SAH collapse rule.
// RANGES: the covered range of sorted positions is not known yet (PLOC topology, leaves already in depth-first order) and is
// derived here from the children's ranges.
template <bool RANGES>
__global__ void __launch_bounds__(256)
k_fit(const int n, const int* __restrict__ vals, const float4* __restrict__ lo, const float4* __restrict__ hi,
      float4* __restrict__ slo, float4* __restrict__ shi, const int* __restrict__ child_l, const int* __restrict__ child_r,
      int* first, int* last, const int* __restrict__ parent_i, const int* __restrict__ parent_p,
      float4* nlo, float4* nhi, float* cost, int* __restrict__ collapse, int* flag, const int max_leaf, const float node_cost)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int prim = vals[p];
    const float4 l0 = lo[prim], h0 = hi[prim];
    slo[p] = l0; shi[p] = h0;
    if (n < 2) return;
    int cur = parent_p[p];
    while (cur >= 0) {
        __threadfence();
        if (atomicAdd(flag + cur, 1) == 0) return;
        const int cl = child_l[cur], cr = child_r[cur];
        float4 al, ah, bl, bh; float ca, cb;
        if (cl < 0) { al = __ldcg(slo + ~cl); ah = __ldcg(shi + ~cl); ca = half_area(al, ah); }
        else { al = __ldcg(nlo + cl); ah = __ldcg(nhi + cl); ca = __ldcg(cost + cl); }
        if (cr < 0) { bl = __ldcg(slo + ~cr); bh = __ldcg(shi + ~cr); cb = half_area(bl, bh); }
        else { bl = __ldcg(nlo + cr); bh = __ldcg(nhi + cr); cb = __ldcg(cost + cr); }
        const float4 l = make_float4(fminf(al.x, bl.x), fminf(al.y, bl.y), fminf(al.z, bl.z), 0.0f);
        const float4 h = make_float4(fmaxf(ah.x, bh.x), fmaxf(ah.y, bh.y), fmaxf(ah.z, bh.z), 0.0f);
        const float a = half_area(l, h);
        if (RANGES) {
            first[cur] = cl < 0 ? ~cl : __ldcg(first + cl);
            last[cur] = cr < 0 ? ~cr : __ldcg(last + cr);
        }
        const int cnt = last[cur] - first[cur] + 1;
        const float leaf_cost = (float)cnt * a, split_cost = node_cost * a + ca + cb;
        const bool col = cnt <= max_leaf && leaf_cost <= split_cost;
        nlo[cur] = l; nhi[cur] = h; cost[cur] = col ? leaf_cost : split_cost; collapse[cur] = col ? 1 : 0;
        cur = parent_i[cur];
    }
}

// One BFS level: order[begin, end) holds tree nodes that become emitted inner nodes; their non-collapsed
// internal children are appended behind (so a child's emitted index is always larger than its parent's).
__global__ void __launch_bounds__(256)
k_bfs_level(int* __restrict__ order, const int begin, const int end, const int* __restrict__ child_l, const int* __restrict__ child_r,
            const int* __restrict__ collapse, const int* __restrict__ first, int* __restrict__ n_emitted, int* __restrict__ emit_id,
            int* __restrict__ leaf_start)
{
    const int i = begin + blockIdx.x * blockDim.x + threadIdx.x;
    const bool on = i < end;
    int c[2] = {0, 0};
    bool inner[2] = {false, false};
    if (on) {
        const int t = order[i];
        c[0] = child_l[t]; c[1] = child_r[t];
        for (int k = 0; k < 2; ++k) {
            if (c[k] < 0) leaf_start[~c[k]] = 1;
            else if (collapse[c[k]]) leaf_start[first[c[k]]] = 1;
            else inner[k] = true;
        }
    }
    // warp-aggregated allocation of emitted indices
    const unsigned lane = threadIdx.x & 31u;
    const int want = (inner[0] ? 1 : 0) + (inner[1] ? 1 : 0);
    int incl = want;
    for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, o); if ((int)lane >= o) incl += v; }
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    int base = 0;
    if (lane == 31 && total) base = atomicAdd(n_emitted, total);
    base = __shfl_sync(0xffffffffu, base, 31);
    int e = base + incl - want;
    for (int k = 0; k < 2; ++k)
        if (inner[k]) { emit_id[c[k]] = e; order[e] = c[k]; ++e; }
}

__global__ void __launch_bounds__(256)
k_emit_nodes(const int* __restrict__ order, const int n_emitted, const int* __restrict__ child_l, const int* __restrict__ child_r,
             const int* __restrict__ collapse, const int* __restrict__ first, const int* __restrict__ emit_id,
             const int* __restrict__ leaf_incl, const float4* __restrict__ slo, const float4* __restrict__ shi,
             const float4* __restrict__ nlo, const float4* __restrict__ nhi, float4* __restrict__ nodes)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_emitted) return;
    const int t = order[e];
    const int c[2] = {child_l[t], child_r[t]};
    float4 l[2], h[2]; int ptr[2];
    for (int k = 0; k < 2; ++k) {
        if (c[k] < 0) {
            const int p = ~c[k];
            l[k] = slo[p]; h[k] = shi[p]; ptr[k] = ~(3 * p + leaf_incl[p] - 1);
        } else {
            l[k] = nlo[c[k]]; h[k] = nhi[c[k]];
            if (collapse[c[k]]) { const int p = first[c[k]]; ptr[k] = ~(3 * p + leaf_incl[p] - 1); }
            else ptr[k] = emit_id[c[k]] * 64;             // byte offset, CudaBVH.cpp:161
        }
    }
    nodes[4 * (size_t)e + 0] = make_float4(l[0].x, h[0].x, l[0].y, h[0].y);      // CudaBVH.cpp:219-222
    nodes[4 * (size_t)e + 1] = make_float4(l[1].x, h[1].x, l[1].y, h[1].y);
    nodes[4 * (size_t)e + 2] = make_float4(l[0].z, h[0].z, l[1].z, h[1].z);
    nodes[4 * (size_t)e + 3] = make_float4(__int_as_float(ptr[0]), __int_as_float(ptr[1]), 0.0f, 0.0f);
}

__global__ void __launch_bounds__(256)
k_emit_tris(const float* __restrict__ verts, const int* __restrict__ tri, const int* __restrict__ vals, const int n,
            const int* __restrict__ leaf_start, const int* __restrict__ leaf_incl, float4* __restrict__ tris, int* __restrict__ indices)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int prim = vals[p];
    const size_t slot = 3 * (size_t)p + (size_t)(leaf_incl[p] - 1);
    const int a = tri[3 * prim], b = tri[3 * prim + 1], c = tri[3 * prim + 2];
    // v0.x == -0.0 has the terminator's bit pattern and would end the leaf early in the reference kernel
    // (cudaUtils.h:413); +0.0 is the same point
    tris[slot + 0] = make_float4(verts[3 * a] + 0.0f, verts[3 * a + 1], verts[3 * a + 2], 0.0f);      // CudaBVH.cpp:288-290
    tris[slot + 1] = make_float4(verts[3 * b], verts[3 * b + 1], verts[3 * b + 2], 0.0f);
    tris[slot + 2] = make_float4(verts[3 * c], verts[3 * c + 1], verts[3 * c + 2], 0.0f);
    indices[slot] = prim; indices[slot + 1] = 0; indices[slot + 2] = 0;                                   // :201-203
    if (p == n - 1 || leaf_start[p + 1]) {
        const float term = __int_as_float((int)0x80000000u);
        tris[slot + 3] = make_float4(term, term, term, term);                                            // :207-208
        indices[slot + 3] = 0;                                                                            // :212
    }
}

// n == 1: createCompact cannot express a root that is a leaf (assert, CudaBVH.cpp:141); both children of the
// root point at the same one-triangle leaf.
__global__ void k_single_root(const float4* __restrict__ slo, const float4* __restrict__ shi, float4* __restrict__ nodes)
{
    const float4 l = slo[0], h = shi[0];
    nodes[0] = make_float4(l.x, h.x, l.y, h.y);
    nodes[1] = make_float4(l.x, h.x, l.y, h.y);
    nodes[2] = make_float4(l.z, h.z, l.z, h.z);
    nodes[3] = make_float4(__int_as_float(~0), __int_as_float(~0), 0.0f, 0.0f);
}

// ---- PLOC topology (Meister & Bittner 2018, "Parallel Locally-Ordered Clustering for Bounding Volume Hierarchy Construction") ----
// Bottom-up agglomerative clustering over the Morton-sorted primitives: every cluster looks for the neighbour within `r`
// array positions whose union box with it has the smallest area; two clusters that chose each other merge into a new node;
// the cluster array is compacted (order kept) and the step repeats until one cluster -- the root -- is left.  Unlike the
// radix tree, the topology follows box AREAS, not bit prefixes of the centres: on scenes with objects of mixed size the SAH
// cost is ~1.7x lower (profiles/r01_tree_quality.jsonl).  Everything is deterministic: ties go to the smaller position, node
// ids and output positions come from prefix sums.
#define PLOC_R_MAX 32

__global__ void __launch_bounds__(256)
k_ploc_init(const int n, const int* __restrict__ vals, const float4* __restrict__ lo, const float4* __restrict__ hi,
            int* __restrict__ cid, float4* __restrict__ blo, float4* __restrict__ bhi)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int prim = vals[i];
    cid[i] = ~i; blo[i] = lo[prim]; bhi[i] = hi[prim];
}

__global__ void __launch_bounds__(256)
k_ploc_nn(const float4* __restrict__ blo, const float4* __restrict__ bhi, const int m, const int r, int* __restrict__ nn)
{
    __shared__ float4 s_lo[256 + 2 * PLOC_R_MAX], s_hi[256 + 2 * PLOC_R_MAX];
    const int base = blockIdx.x * 256 - r;
    for (int k = threadIdx.x; k < 256 + 2 * r; k += 256) {
        const int g = base + k;
        if (g >= 0 && g < m) { s_lo[k] = blo[g]; s_hi[k] = bhi[g]; }
    }
    __syncthreads();
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= m) return;
    const float4 l = s_lo[threadIdx.x + r], h = s_hi[threadIdx.x + r];
    const int j0 = i - r > 0 ? i - r : 0, j1 = i + r < m - 1 ? i + r : m - 1;
    float best = 3.4e38f; int bj = -1;
    for (int j = j0; j <= j1; ++j) {
        if (j == i) continue;
        const float4 a = s_lo[j - base], b = s_hi[j - base];
        const float dx = fmaxf(h.x, b.x) - fminf(l.x, a.x), dy = fmaxf(h.y, b.y) - fminf(l.y, a.y), dz = fmaxf(h.z, b.z) - fminf(l.z, a.z);
        const float area = dx * dy + dy * dz + dz * dx;
        if (area < best) { best = area; bj = j; }          // ascending j, strict <: ties go to the smaller position
    }
    if (bj < 0) bj = i > j0 ? i - 1 : i + 1;                 // areas overflowed: any neighbour
    nn[i] = bj;
}

// per cluster: (1 << 32) if it survives the step (not the higher half of a merging pair) + 1 if it leads a merge
__global__ void __launch_bounds__(256)
k_ploc_flags(const int* __restrict__ nn, const int m, unsigned long long* __restrict__ fl)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const int j = nn[i];
    const bool mutual = nn[j] == i;
    fl[i] = (mutual && i > j) ? 0ull : ((1ull << 32) | (mutual ? 1ull : 0ull));
}

__global__ void __launch_bounds__(256)
k_ploc_merge(const int* __restrict__ nn, const unsigned long long* __restrict__ incl, const int m, const int id_top,
             const int* __restrict__ cid_in, const float4* __restrict__ blo_in, const float4* __restrict__ bhi_in,
             int* __restrict__ cid_out, float4* __restrict__ blo_out, float4* __restrict__ bhi_out,
             int* __restrict__ child_l, int* __restrict__ child_r, int* __restrict__ parent_i, int* __restrict__ parent_p,
             int* __restrict__ count)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const int j = nn[i];
    const bool mutual = nn[j] == i;
    if (mutual && i > j) return;
    const unsigned long long s = incl[i];
    const int pos = (int)(s >> 32) - 1;
    float4 l = blo_in[i], h = bhi_in[i];
    int id = cid_in[i];
    if (mutual) {
        const int a = id, b = cid_in[j];
        id = id_top - ((int)(s & 0xffffffffull) - 1);        // later nodes get smaller ids: the root ends up as node 0
        child_l[id] = a; child_r[id] = b;
        if (a < 0) parent_p[~a] = id; else parent_i[a] = id;
        if (b < 0) parent_p[~b] = id; else parent_i[b] = id;
        count[id] = (a < 0 ? 1 : count[a]) + (b < 0 ? 1 : count[b]);
        const float4 l2 = blo_in[j], h2 = bhi_in[j];
        l = make_float4(fminf(l.x, l2.x), fminf(l.y, l2.y), fminf(l.z, l2.z), 0.0f);
        h = make_float4(fmaxf(h.x, h2.x), fmaxf(h.y, h2.y), fmaxf(h.z, h2.z), 0.0f);
    }
    cid_out[pos] = id; blo_out[pos] = l; bhi_out[pos] = h;
}

// depth-first position of every leaf: the number of leaves left of it = sum, over the ancestors it hangs under on the right,
// of the left sibling's leaf count
__global__ void __launch_bounds__(256)
k_ploc_positions(const int n, const int* __restrict__ child_l, const int* __restrict__ child_r, const int* __restrict__ parent_i,
                 const int* __restrict__ parent_p, const int* __restrict__ count, int* __restrict__ posmap, int* __restrict__ too_deep)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    int pos = 0, child = ~p, cur = parent_p[p], steps = 0;
    while (cur >= 0) {
        if (child_r[cur] == child) { const int l = child_l[cur]; pos += l < 0 ? 1 : count[l]; }
        child = cur; cur = parent_i[cur];
        if (++steps > 100000) { atomicExch(too_deep, 4); break; }
    }
    posmap[p] = pos;
}

__global__ void __launch_bounds__(256)
k_ploc_remap(const int n, const int* __restrict__ posmap, const int* __restrict__ vals_in, const int* __restrict__ parent_p_in,
             int* __restrict__ vals_out, int* __restrict__ parent_p_out, int* __restrict__ child_l, int* __restrict__ child_r)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n) { const int q = posmap[p]; vals_out[q] = vals_in[p]; parent_p_out[q] = parent_p_in[p]; }
    if (p < n - 1) {
        const int a = child_l[p], b = child_r[p];
        if (a < 0) child_l[p] = ~posmap[~a];
        if (b < 0) child_r[p] = ~posmap[~b];
    }
}

} // namespace

cudaError_t gptk_build_lbvh(const float* verts, size_t n_verts, const int* tri, size_t n_tris, int max_leaf, int ploc_radius,
                            BuildResult* out, cudaStream_t st)
{
    *out = BuildResult{};
    if (n_tris == 0) return cudaSuccess;
    if (n_tris > 0x1fffffffull / 3 || n_verts > 0x7fffffffull) { out->status = 3; return cudaSuccess; }
    const int n = (int)n_tris;
    if (max_leaf < 1) max_leaf = 1;
    if (max_leaf > 64) max_leaf = 64;
    if (ploc_radius > PLOC_R_MAX) ploc_radius = PLOC_R_MAX;
    if (n < 2 || ploc_radius < 0) ploc_radius = 0;
    static const float node_cost = []{ const char* v = getenv("GPT_SAH_NODE_COST"); return (v && *v) ? (float)atof(v) : 1.0f; }();
    cudaError_t e = cudaSuccess;
    // scratch
    float4 *lo = nullptr, *hi = nullptr, *slo = nullptr, *shi = nullptr, *nlo = nullptr, *nhi = nullptr;
    unsigned long long *keys = nullptr, *keys2 = nullptr;
    int *vals = nullptr, *vals2 = nullptr, *ibuf = nullptr, *small = nullptr;
    float* cost = nullptr;
    float4* pbox = nullptr; int* pint = nullptr; unsigned long long* pfl = nullptr;      // PLOC scratch
    void* tmp = nullptr; size_t tmp_bytes = 0, tmp2 = 0;
    float4* nodes = nullptr; float4* tris_out = nullptr; int* idx_out = nullptr;
    const int g = (n + 255) / 256;
    const size_t ni = (size_t)(n > 1 ? n - 1 : 1);
#define CK(x) do { e = (x); if (e != cudaSuccess) goto fail; } while (0)
    CK(cudaMalloc(&lo, sizeof(float4) * n)); CK(cudaMalloc(&hi, sizeof(float4) * n));
    CK(cudaMalloc(&slo, sizeof(float4) * n)); CK(cudaMalloc(&shi, sizeof(float4) * n));
    CK(cudaMalloc(&nlo, sizeof(float4) * ni)); CK(cudaMalloc(&nhi, sizeof(float4) * ni));
    CK(cudaMalloc(&cost, sizeof(float) * ni));
    CK(cudaMalloc(&keys, sizeof(unsigned long long) * n)); CK(cudaMalloc(&keys2, sizeof(unsigned long long) * n));
    CK(cudaMalloc(&vals, sizeof(int) * n)); CK(cudaMalloc(&vals2, sizeof(int) * n));
    // int arrays: child_l child_r first last parent_i collapse flag emit_id order (ni each) + parent_p leaf_start leaf_incl (n each)
    CK(cudaMalloc(&ibuf, sizeof(int) * (9 * ni + 3 * (size_t)n)));
    CK(cudaMalloc(&small, sizeof(int) * 8));
    {
        int* child_l = ibuf; int* child_r = child_l + ni; int* first = child_r + ni; int* last = first + ni;
        int* parent_i = last + ni; int* collapse = parent_i + ni; int* flag = collapse + ni; int* emit_id = flag + ni;
        int* order = emit_id + ni; int* parent_p = order + ni; int* leaf_start = parent_p + n; int* leaf_incl = leaf_start + n;
        int* scene = small; int* bad = small + 6; int* n_emitted = small + 7;
        const int init[8] = {0x7f7fffff, 0x7f7fffff, 0x7f7fffff, (int)0x80800000, (int)0x80800000, (int)0x80800000, 0, 1};
        CK(cudaMemcpyAsync(small, init, sizeof(init), cudaMemcpyHostToDevice, st));
        CK(cudaMemsetAsync(flag, 0, sizeof(int) * ni, st));
        CK(cudaMemsetAsync(collapse, 0, sizeof(int) * ni, st));
        CK(cudaMemsetAsync(leaf_start, 0, sizeof(int) * (size_t)n, st));
        k_prim_bounds<<<g, 256, 0, st>>>(verts, (int)n_verts, tri, n, lo, hi, scene, bad);
        k_morton<<<g, 256, 0, st>>>(lo, hi, n, scene, keys, vals);
        CK(cudaGetLastError());
        CK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys, keys2, vals, vals2, n, 0, 63, st));
        CK(cub::DeviceScan::InclusiveSum(nullptr, tmp2, leaf_start, leaf_incl, n, st));
        if (tmp2 > tmp_bytes) tmp_bytes = tmp2;
        CK(cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 16));
        CK(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys, keys2, vals, vals2, n, 0, 63, st));
        int h_small[8];
        int n_em = 1, depth = 1;
        const int* svals = vals2;            // primitive at each sorted position (PLOC: re-ordered depth first below)
        int ploc_steps = 0;
        if (ploc_radius > 0) {
            // cluster arrays, double-buffered: ids, boxes; nearest neighbour; flags + their prefix sums; leaf counts; position map
            CK(cudaMalloc(&pbox, sizeof(float4) * 4 * (size_t)n));
            CK(cudaMalloc(&pint, sizeof(int) * 6 * (size_t)n));
            CK(cudaMalloc(&pfl, sizeof(unsigned long long) * 2 * (size_t)n));
            size_t tmp3 = 0;
            CK(cub::DeviceScan::InclusiveSum(nullptr, tmp3, pfl, pfl + n, n, st));
            if (tmp3 > tmp_bytes) { cudaFree(tmp); tmp = nullptr; tmp_bytes = tmp3; CK(cudaMalloc(&tmp, tmp_bytes)); }
            float4* blo[2] = {pbox, pbox + 2 * (size_t)n}; float4* bhi[2] = {pbox + n, pbox + 3 * (size_t)n};
            int* cid[2] = {pint, pint + n}; int* nn = pint + 2 * (size_t)n; int* count = pint + 3 * (size_t)n;
            int* posmap = pint + 4 * (size_t)n; int* parent_p0 = pint + 5 * (size_t)n;
            CK(cudaMemcpyAsync(h_small, small, sizeof(h_small), cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            if (h_small[6]) { out->status = h_small[6]; goto done; }          // non-finite input: no area ordering
            k_ploc_init<<<g, 256, 0, st>>>(n, vals2, lo, hi, cid[0], blo[0], bhi[0]);
            int m = n, made = 0, cur = 0;
            while (m > 1) {
                const int gm = (m + 255) / 256;
                k_ploc_nn<<<gm, 256, 0, st>>>(blo[cur], bhi[cur], m, ploc_radius, nn);
                k_ploc_flags<<<gm, 256, 0, st>>>(nn, m, pfl);
                CK(cub::DeviceScan::InclusiveSum(tmp, tmp_bytes, pfl, pfl + n, m, st));
                k_ploc_merge<<<gm, 256, 0, st>>>(nn, pfl + n, m, n - 2 - made, cid[cur], blo[cur], bhi[cur], cid[cur ^ 1], blo[cur ^ 1], bhi[cur ^ 1],
                                                 child_l, child_r, parent_i, parent_p0, count);
                CK(cudaGetLastError());
                unsigned long long tot = 0;
                CK(cudaMemcpyAsync(&tot, pfl + n + (m - 1), sizeof(tot), cudaMemcpyDeviceToHost, st));
                CK(cudaStreamSynchronize(st));
                const int m2 = (int)(tot >> 32), created = (int)(tot & 0xffffffffull);
                if (created <= 0 || m2 != m - created || ++ploc_steps > 2048) { out->status = 5; goto done; }
                made += created; m = m2; cur ^= 1;
            }
            if (made != n - 1) { out->status = 5; goto done; }
            const int minus1 = -1;
            CK(cudaMemcpyAsync(parent_i, &minus1, sizeof(int), cudaMemcpyHostToDevice, st));
            k_ploc_positions<<<g, 256, 0, st>>>(n, child_l, child_r, parent_i, parent_p0, count, posmap, bad);
            k_ploc_remap<<<g, 256, 0, st>>>(n, posmap, vals2, parent_p0, vals, parent_p, child_l, child_r);
            svals = vals;
            k_fit<true><<<g, 256, 0, st>>>(n, svals, lo, hi, slo, shi, child_l, child_r, first, last, parent_i, parent_p, nlo, nhi, cost, collapse, flag, max_leaf, node_cost);
            CK(cudaGetLastError());
        } else if (n >= 2) {
            k_radix_tree<<<g, 256, 0, st>>>(keys2, n, child_l, child_r, first, last, parent_i, parent_p);
            k_fit<false><<<g, 256, 0, st>>>(n, svals, lo, hi, slo, shi, child_l, child_r, first, last, parent_i, parent_p, nlo, nhi, cost, collapse, flag, max_leaf, node_cost);
            CK(cudaGetLastError());
        }
        if (n >= 2) {
            // the root is always emitted (createCompact needs two children)
            const int zero = 0;
            CK(cudaMemcpyAsync(order, &zero, sizeof(int), cudaMemcpyHostToDevice, st));
            CK(cudaMemcpyAsync(emit_id, &zero, sizeof(int), cudaMemcpyHostToDevice, st));
            int begin = 0, end = 1;
            while (begin < end) {
                k_bfs_level<<<(end - begin + 255) / 256, 256, 0, st>>>(order, begin, end, child_l, child_r, collapse, first, n_emitted, emit_id, leaf_start);
                CK(cudaMemcpyAsync(h_small, small, sizeof(h_small), cudaMemcpyDeviceToHost, st));
                CK(cudaStreamSynchronize(st));
                if (h_small[6]) { out->status = h_small[6]; goto done; }
                begin = end; end = h_small[7];
                if (begin < end) ++depth;
                if (depth > 4096) { out->status = 4; goto done; }
            }
            n_em = end;
        } else {
            k_fit<false><<<g, 256, 0, st>>>(n, svals, lo, hi, slo, shi, child_l, child_r, first, last, parent_i, parent_p, nlo, nhi, cost, collapse, flag, max_leaf, node_cost);
            const int one = 1;
            CK(cudaMemcpyAsync(leaf_start, &one, sizeof(int), cudaMemcpyHostToDevice, st));
            CK(cudaMemcpyAsync(h_small, small, sizeof(h_small), cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            if (h_small[6]) { out->status = h_small[6]; goto done; }
        }
        if ((size_t)n_em * 64 > 0x7fffffffull) { out->status = 3; goto done; }          // child byte offsets are ints
        CK(cub::DeviceScan::InclusiveSum(tmp, tmp_bytes, leaf_start, leaf_incl, n, st));
        int n_leaves = 0;
        CK(cudaMemcpyAsync(&n_leaves, leaf_incl + (n - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        const size_t n_slots = 3 * (size_t)n + (size_t)n_leaves;
        if (n_slots > 0x7fffffffull) { out->status = 3; goto done; }
        CK(cudaMalloc(&nodes, sizeof(float4) * 4 * (size_t)n_em));
        CK(cudaMalloc(&tris_out, sizeof(float4) * n_slots));
        CK(cudaMalloc(&idx_out, sizeof(int) * n_slots));
        if (n >= 2)
            k_emit_nodes<<<(n_em + 255) / 256, 256, 0, st>>>(order, n_em, child_l, child_r, collapse, first, emit_id, leaf_incl, slo, shi, nlo, nhi, nodes);
        else
            k_single_root<<<1, 1, 0, st>>>(slo, shi, nodes);
        k_emit_tris<<<g, 256, 0, st>>>(verts, tri, svals, n, leaf_start, leaf_incl, tris_out, idx_out);
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(st));
        out->nodes = nodes; out->n_node_vec4 = 4 * (size_t)n_em;
        out->tris = tris_out; out->n_tri_vec4 = n_slots;
        out->indices = idx_out; out->n_indices = n_slots;
        out->n_leaves = n_leaves; out->depth = depth; out->launches = 6 + depth + (ploc_radius > 0 ? 4 + 4 * ploc_steps : 0);
        nodes = nullptr; tris_out = nullptr; idx_out = nullptr;
    }
done:
fail:
    cudaFree(lo); cudaFree(hi); cudaFree(slo); cudaFree(shi); cudaFree(nlo); cudaFree(nhi); cudaFree(cost);
    cudaFree(keys); cudaFree(keys2); cudaFree(vals); cudaFree(vals2); cudaFree(ibuf); cudaFree(small); cudaFree(tmp);
    cudaFree(nodes); cudaFree(tris_out); cudaFree(idx_out); cudaFree(pbox); cudaFree(pint); cudaFree(pfl);
    return e;
#undef CK
}
