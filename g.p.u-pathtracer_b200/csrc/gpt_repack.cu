// gpt_repack.cu -- device-side re-pack of the arrays CudaBVH::createCompact emits (CudaBVH.cpp:121-270)
// into the layout described in gpt_scene.cuh.  The SAME binary tree, child order and per-leaf triangle
// order are kept, so traversal order -- and therefore tie-breaking -- is unchanged.
//
// Three passes over the 2*n_inner child slots:
//   1. count : a leaf child (~slot, CudaBVH.cpp:177) is walked exactly the way the reference kernel reads
//              it (v0 first; bits(v0.x)==0x80000000 ends the leaf, cudaUtils.h:405-416) -> triangles per leaf
//   2. scan  : exclusive prefix sum (CUB) -> first packed triangle of each leaf
//   3. write : packed triangles (v0, e1, e2 + index / slot / triangles left in the leaf) and patched child pointers
#include "gpt_kernels.h"
#include <cub/device/device_scan.cuh>

#define LEAF_TERM 0x80000000u

__global__ void k_count_leaf_tris(const float4* __restrict__ ref_nodes, int n_inner, const float4* __restrict__ ref_tris,
                                  size_t n_tri_vec4, size_t n_node_vec4, int* __restrict__ counts, int* __restrict__ bad)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 2 * n_inner) return;
    const int4 ch = *reinterpret_cast<const int4*>(ref_nodes + 4 * (i >> 1) + 3);
    const int c = (i & 1) ? ch.y : ch.x;
    int cnt = 0;
    if (c < 0) {
        size_t slot = (size_t)(~c);
        while (true) {
            if (slot >= n_tri_vec4) { atomicExch(bad, 1); break; }              // ran off the array
            if (__float_as_uint(ref_tris[slot].x) == LEAF_TERM) break;
            if (slot + 2 >= n_tri_vec4) { atomicExch(bad, 1); break; }
            ++cnt; slot += 3;
        }
    } else {
        // inner child: byte offset, 64-B aligned, inside the node array (CudaBVH.cpp:161)
        // and, as createCompact appends children after their parent, strictly behind it (rules out cycles)
        if ((c & 63) != 0 || (size_t)(c >> 4) + 4 > n_node_vec4 || (c >> 6) <= (i >> 1)) atomicExch(bad, 3);
    }
    counts[i] = cnt;
}

__global__ void k_write_packed(const float4* __restrict__ ref_nodes, int n_inner, const float4* __restrict__ ref_tris,
                               const int* __restrict__ ref_indices, size_t n_indices, const int* __restrict__ offsets,
                               const int n_total, float4* __restrict__ nodes, float4* __restrict__ tris)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 2 * n_inner) return;
    const int node = i >> 1;
    const int4 ch = *reinterpret_cast<const int4*>(ref_nodes + 4 * node + 3);
    const int c = (i & 1) ? ch.y : ch.x;
    int packed_child;
    if (c >= 0) {
        packed_child = c >> 4;                       // byte offset -> float4 index
    } else {
        const int base = offsets[i];
        // A leaf whose FIRST v0.x is -0.0 reads as empty in the reference kernel (cudaUtils.h:413, the
        // "-0.0 hazard"); it points at the shared null triangle appended behind the last real one.
        int slot = ~c, k = 0;
        const int cnt = offsets[i + 1] - base;                  // triangles in this leaf (pass 1)
        packed_child = (__float_as_uint(ref_tris[~c].x) == LEAF_TERM) ? ~n_total : ~base;
        while (__float_as_uint(ref_tris[slot].x) != LEAF_TERM) {
            const float4 v0 = ref_tris[slot], v1 = ref_tris[slot + 1], v2 = ref_tris[slot + 2];
            const int orig = ((size_t)slot < n_indices) ? ref_indices[slot] : -1;
            float4* dst = tris + GPT_TRI_STRIDE * (size_t)(base + k);
            dst[0] = make_float4(v0.x, v0.y, v0.z, __int_as_float(orig));
            dst[GPT_TRI_B] = make_float4(v1.x - v0.x, v1.y - v0.y, v1.z - v0.z, __int_as_float(slot));
            dst[GPT_TRI_B + 1] = make_float4(v2.x - v0.x, v2.y - v0.y, v2.z - v0.z, __int_as_float(cnt - 1 - k));

            ++k; slot += 3;
        }
    }
    // boxes are copied by the even lane of the pair; child pointers written per child
    if ((i & 1) == 0) {
        nodes[4 * node + 0] = ref_nodes[4 * node + 0];
        nodes[4 * node + 1] = ref_nodes[4 * node + 1];
        nodes[4 * node + 2] = ref_nodes[4 * node + 2];
    }
    reinterpret_cast<int*>(nodes + 4 * node + 3)[i & 1] = packed_child;
    if ((i & 1) == 0) { reinterpret_cast<int*>(nodes + 4 * node + 3)[2] = 0; reinterpret_cast<int*>(nodes + 4 * node + 3)[3] = 0; }
    if (i == 0) {   // null triangle: zero area -> det == 0 -> never hit; 0 triangles follow it in its leaf
        float4* nt = tris + GPT_TRI_STRIDE * (size_t)n_total;
        nt[0] = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
        nt[GPT_TRI_B] = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
        nt[GPT_TRI_B + 1] = make_float4(0.f, 0.f, 0.f, __int_as_float(0));

    }
}

// Depth of the packed tree (root = level 1), level by level: a node's level is its parent's + 1, children sit behind their
// parents (checked by k_count_leaf_tris), so pass k fixes every node of level k + 1.  The traversal keeps one stack entry per
// inner level below the root plus the sentinel (cudaUtils.h:22,259 sizes it 64 for the builder's MaxDepth 64,
// SplitBVHBuilder.hpp:15): a tree of GPT_STACK_SIZE or more levels would write past the per-thread stack, so it is rejected.
__global__ void k_node_levels(const float4* __restrict__ nodes, const int n_inner, int* __restrict__ level, const int pass, int* __restrict__ changed)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_inner || level[i] != pass) return;
    const int4 ch = *reinterpret_cast<const int4*>(nodes + 4 * (size_t)i + 3);
    if (ch.x >= 0) { level[ch.x >> 2] = pass + 1; *changed = 1; }
    if (ch.y >= 0) { level[ch.y >> 2] = pass + 1; *changed = 1; }
}

cudaError_t gptk_repack(const float4* ref_nodes, size_t n_node_vec4, const float4* ref_tris, size_t n_tri_vec4,
                        const int* ref_indices, size_t n_indices, RepackResult* out, cudaStream_t st)
{
    out->nodes = nullptr; out->tris = nullptr; out->n_nodes = 0; out->n_tris = 0; out->bad_layout = 0; out->depth = 0;
    const int n_inner = (int)(n_node_vec4 / 4);
    if (n_inner == 0) return cudaSuccess;
    if (n_node_vec4 % 4 != 0 || n_node_vec4 / 4 > 0x1fffffffull) { out->bad_layout = 4; return cudaSuccess; }
    const int n_child = 2 * n_inner;
    int *counts = nullptr, *offsets = nullptr, *bad = nullptr;
    void* tmp = nullptr; size_t tmp_bytes = 0;
    cudaError_t e;
#define CK(x) do { e = (x); if (e != cudaSuccess) goto fail; } while (0)
    CK(cudaMalloc(&counts, sizeof(int) * (size_t)(n_child + 1)));
    CK(cudaMalloc(&offsets, sizeof(int) * (size_t)(n_child + 1)));
    CK(cudaMalloc(&bad, sizeof(int)));
    CK(cudaMemsetAsync(bad, 0, sizeof(int), st));
    CK(cudaMemsetAsync(counts, 0, sizeof(int) * (size_t)(n_child + 1), st));
    k_count_leaf_tris<<<(n_child + 127) / 128, 128, 0, st>>>(ref_nodes, n_inner, ref_tris, n_tri_vec4, n_node_vec4, counts, bad);
    CK(cudaGetLastError());
    CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, counts, offsets, n_child + 1, st));
    CK(cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 16));
    CK(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, counts, offsets, n_child + 1, st));
    {
        int h_bad = 0, h_total = 0;
        CK(cudaMemcpyAsync(&h_bad, bad, sizeof(int), cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(&h_total, offsets + n_child, sizeof(int), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        if (h_bad) { out->bad_layout = h_bad; goto done; }
        CK(cudaMalloc(&out->nodes, sizeof(float4) * 4 * (size_t)n_inner));
        CK(cudaMalloc(&out->tris, sizeof(float4) * GPT_TRI_STRIDE * ((size_t)h_total + 1)));
        k_write_packed<<<(n_child + 127) / 128, 128, 0, st>>>(ref_nodes, n_inner, ref_tris, ref_indices, n_indices, offsets, h_total, out->nodes, out->tris);
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(st));
        out->n_nodes = n_inner; out->n_tris = h_total;
        // tree depth (the counts array is free again: reuse it for the levels)
        CK(cudaMemsetAsync(counts, 0, sizeof(int) * (size_t)n_inner, st));
        { const int one = 1; CK(cudaMemcpyAsync(counts, &one, sizeof(int), cudaMemcpyHostToDevice, st)); }
        int depth = 1;
        for (; depth <= GPT_STACK_SIZE; ++depth) {
            int h_changed = 0;
            CK(cudaMemsetAsync(bad, 0, sizeof(int), st));
            k_node_levels<<<(n_inner + 255) / 256, 256, 0, st>>>(out->nodes, n_inner, counts, depth, bad);
            CK(cudaGetLastError());
            CK(cudaMemcpyAsync(&h_changed, bad, sizeof(int), cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            if (!h_changed) break;
        }
        out->depth = depth;
        if (depth >= GPT_STACK_SIZE) {
            cudaFree(out->nodes); cudaFree(out->tris); out->nodes = nullptr; out->tris = nullptr; out->n_nodes = out->n_tris = 0;
            out->bad_layout = 5;
        }
    }
done:
    cudaFree(counts); cudaFree(offsets); cudaFree(bad); cudaFree(tmp);
    return cudaSuccess;
fail:
    cudaFree(counts); cudaFree(offsets); cudaFree(bad); cudaFree(tmp);
    if (out->nodes) cudaFree(out->nodes);
    if (out->tris) cudaFree(out->tris);
    out->nodes = nullptr; out->tris = nullptr;
    return e;
#undef CK
}

// ---- a second tree over the SAME triangle references, built by this library's own builder (gpt_build.cu) ----------------------
// The reference's host SBVH is tuned for its own kernel (leaves of up to 8 triangles: 6.5 triangle tests per ray on the bench
// scene); the while-while extend kernel here is faster on the PLOC tree with SAH-collapsed leaves (+17 % rays/s on the bench scene,
// +18 % on the 100 k-triangle dragon).  gptk_rebuild makes that tree from what the caller uploaded: every triangle reference the
// re-pack kept (duplicates of spatial splits included -- the same triangle twice gives the same t and the same face index) goes
// into the builder as a triangle soup with the caller's exact vertices, the builder's arrays go through the same re-pack, and
// the face index / caller slot of each packed triangle are restored from the first packing.  Packed rows (v0, v1 - v0, v2 - v0)
// are bit-identical to the first packing's, so every ray-triangle distance is; only the ORDER in which triangles are met differs,
// i.e. which of two triangles with bit-equal t wins (the 4-wide tree's caveat).
__global__ void k_soup(const float4* __restrict__ packed, const int n, const float4* __restrict__ ref_tris, float* __restrict__ verts, int* __restrict__ faces)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int slot = __float_as_int(packed[GPT_TRI_STRIDE * (size_t)i + GPT_TRI_B].w);
    for (int k = 0; k < 3; ++k) {
        const float4 v = ref_tris[(size_t)slot + k];
        verts[9 * (size_t)i + 3 * k] = v.x; verts[9 * (size_t)i + 3 * k + 1] = v.y; verts[9 * (size_t)i + 3 * k + 2] = v.z;
        faces[3 * (size_t)i + k] = 3 * i + k;
    }
}

__global__ void k_restore_ids(float4* __restrict__ alt, const int n_alt, const float4* __restrict__ packed, const int n, int* __restrict__ bad)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_alt) return;
    const int i = __float_as_int(alt[GPT_TRI_STRIDE * (size_t)j].w);             // soup triangle = index in the first packing
    if (i < 0 || i >= n) { atomicExch(bad, 1); return; }
    const float4 a = packed[GPT_TRI_STRIDE * (size_t)i], b = packed[GPT_TRI_STRIDE * (size_t)i + GPT_TRI_B];
    const float4 ra = alt[GPT_TRI_STRIDE * (size_t)j], rb = alt[GPT_TRI_STRIDE * (size_t)j + GPT_TRI_B];
    const float4 c = packed[GPT_TRI_STRIDE * (size_t)i + GPT_TRI_B + 1], rc = alt[GPT_TRI_STRIDE * (size_t)j + GPT_TRI_B + 1];
    // the rows must be the first packing's, bit for bit: that is what makes the two trees give the same distances
    if (__float_as_uint(a.x) != __float_as_uint(ra.x) || __float_as_uint(a.y) != __float_as_uint(ra.y) || __float_as_uint(a.z) != __float_as_uint(ra.z) ||
        __float_as_uint(b.x) != __float_as_uint(rb.x) || __float_as_uint(b.y) != __float_as_uint(rb.y) || __float_as_uint(b.z) != __float_as_uint(rb.z) ||
        __float_as_uint(c.x) != __float_as_uint(rc.x) || __float_as_uint(c.y) != __float_as_uint(rc.y) || __float_as_uint(c.z) != __float_as_uint(rc.z)) {
        atomicExch(bad, 2); return;
    }
    alt[GPT_TRI_STRIDE * (size_t)j].w = a.w;                  // original face index
    alt[GPT_TRI_STRIDE * (size_t)j + GPT_TRI_B].w = b.w;      // slot in the caller's triangle array
}

// status: 0 ok (out filled), 1 the builder refused or the tree is too deep, 2 triangles lost or rows differ (internal), 3 too few triangles
cudaError_t gptk_rebuild(const float4* packed_tris, int n_tris, const float4* ref_tris, int max_leaf, int ploc_radius,
                         RepackResult* out, int* status, float* build_ms, unsigned long long* launches, cudaStream_t st)
{
    *out = RepackResult{}; *status = 3; *build_ms = 0.0f;
    if (n_tris < 2) return cudaSuccess;
    float* verts = nullptr; int* faces = nullptr; int* bad = nullptr;
    BuildResult b = BuildResult{};
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    cudaError_t e;
    int h_bad = 0;
#define CK(x) do { e = (x); if (e != cudaSuccess) goto fail; } while (0)
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    CK(cudaEventRecord(e0, st));
    CK(cudaMalloc(&verts, sizeof(float) * 9 * (size_t)n_tris));
    CK(cudaMalloc(&faces, sizeof(int) * 3 * (size_t)n_tris));
    CK(cudaMalloc(&bad, sizeof(int)));
    CK(cudaMemsetAsync(bad, 0, sizeof(int), st));
    k_soup<<<(n_tris + 255) / 256, 256, 0, st>>>(packed_tris, n_tris, ref_tris, verts, faces);
    CK(cudaGetLastError());
    CK(gptk_build_lbvh(verts, 3 * (size_t)n_tris, faces, (size_t)n_tris, max_leaf, ploc_radius, &b, st));
    *launches += 1ull + (unsigned long long)b.launches;
    if (b.status != 0 || !b.nodes) { *status = 1; goto done; }
    CK(gptk_repack(b.nodes, b.n_node_vec4, b.tris, b.n_tri_vec4, b.indices, b.n_indices, out, st));
    *launches += 2;
    if (out->bad_layout || !out->nodes) { *status = 1; goto done; }
    if (out->n_tris != n_tris) { *status = 2; goto done; }
    k_restore_ids<<<(n_tris + 255) / 256, 256, 0, st>>>(out->tris, n_tris, packed_tris, n_tris, bad);
    CK(cudaGetLastError());
    *launches += 1;
    CK(cudaMemcpyAsync(&h_bad, bad, sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaEventRecord(e1, st));
    CK(cudaStreamSynchronize(st));
    cudaEventElapsedTime(build_ms, e0, e1);
    *status = h_bad ? 2 : 0;
done:
    if (*status != 0) { cudaFree(out->nodes); cudaFree(out->tris); *out = RepackResult{}; }
    cudaFree(verts); cudaFree(faces); cudaFree(bad); cudaFree(b.nodes); cudaFree(b.tris); cudaFree(b.indices);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return cudaSuccess;
fail:
    cudaFree(out->nodes); cudaFree(out->tris); *out = RepackResult{};
    cudaFree(verts); cudaFree(faces); cudaFree(bad); cudaFree(b.nodes); cudaFree(b.tris); cudaFree(b.indices);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    *status = 1;
    return e;
#undef CK
}

// ---- the uploaded tree with its leaves split (gptk_split_leaves) ------------------------------------------------------------------
// The reference's SplitBVHBuilder stops at leaves of up to 8 references where its SAH (node cost = triangle cost) says so; the
// kernels here pay 87 instructions per triangle test against 75 per node visit and no longer want them (profiles/r02_schedules.md
// #68: the reference builder run with a leaf cap of 2 gives a tree the extend kernel walks 27 % faster on the dragon -- and 13 %
// faster than the PLOC tree of gptk_rebuild, whose upper levels are not SBVH quality).  A top-down builder's upper levels do not
// depend on where it stops, so that tree is THIS tree with every larger leaf replaced by a small SAH subtree -- which one thread
// per leaf can build: <= 16 references, object splits only, sweep over the three axes.  The packed triangle rows are only permuted
// inside their leaf's range (and their "triangles that follow" lanes rewritten), new inner nodes are appended behind the old ones.
#define SPLIT_MAXT 16

__device__ __forceinline__ float box_half_area(const float3 lo, const float3 hi)
{
    const float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
    return dx * dy + dy * dz + dz * dx;
}

__global__ void k_split_count(const float4* __restrict__ nodes, const int n_inner, const float4* __restrict__ tris, const int max_leaf, int* __restrict__ counts)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 2 * n_inner) return;
    const int4 ch = *reinterpret_cast<const int4*>(nodes + 4 * (size_t)(i >> 1) + 3);
    const int c = (i & 1) ? ch.y : ch.x;
    int extra = 0;
    if (c < 0) {
        const int cnt = __float_as_int(tris[GPT_TRI_STRIDE * (size_t)(~c) + GPT_TRI_B + 1].w) + 1;
        if (cnt > max_leaf && cnt <= SPLIT_MAXT) extra = cnt - 1;          // a binary tree over cnt leaves has at most cnt - 1 inner nodes
    }
    counts[i] = extra;
}

__global__ void __launch_bounds__(64)
k_split_build(const float4* __restrict__ nodes_old, const int n_inner, const float4* __restrict__ tris_old, const float4* __restrict__ ref_tris,
              const int* __restrict__ offsets, const int max_leaf, float4* __restrict__ nodes_new, float4* __restrict__ tris_new)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 2 * n_inner) return;
    if (offsets[i + 1] == offsets[i]) return;                               // not a leaf to split
    const int4 ch = *reinterpret_cast<const int4*>(nodes_old + 4 * (size_t)(i >> 1) + 3);
    const int base = ~((i & 1) ? ch.y : ch.x);
    const int cnt = __float_as_int(tris_old[GPT_TRI_STRIDE * (size_t)base + GPT_TRI_B + 1].w) + 1;
    float3 lo[SPLIT_MAXT], hi[SPLIT_MAXT];
    float right_area[SPLIT_MAXT];
    int perm[SPLIT_MAXT], follow[SPLIT_MAXT];
    for (int k = 0; k < cnt; ++k) {
        const float4 a = tris_old[GPT_TRI_STRIDE * (size_t)(base + k)], b = tris_old[GPT_TRI_STRIDE * (size_t)(base + k) + GPT_TRI_B],
                     c = tris_old[GPT_TRI_STRIDE * (size_t)(base + k) + GPT_TRI_B + 1];
        const int slot = __float_as_int(b.w);
        const float4 v1 = ref_tris[(size_t)slot + 1], v2 = ref_tris[(size_t)slot + 2];      // the caller's exact vertices ...
        const float w1x = a.x + b.x, w1y = a.y + b.y, w1z = a.z + b.z, w2x = a.x + c.x, w2y = a.y + c.y, w2z = a.z + c.z;   // ... and the ones the test sees
        lo[k] = make_float3(fminf(fminf(a.x, fminf(v1.x, v2.x)), fminf(w1x, w2x)), fminf(fminf(a.y, fminf(v1.y, v2.y)), fminf(w1y, w2y)), fminf(fminf(a.z, fminf(v1.z, v2.z)), fminf(w1z, w2z)));
        hi[k] = make_float3(fmaxf(fmaxf(a.x, fmaxf(v1.x, v2.x)), fmaxf(w1x, w2x)), fmaxf(fmaxf(a.y, fmaxf(v1.y, v2.y)), fmaxf(w1y, w2y)), fmaxf(fmaxf(a.z, fmaxf(v1.z, v2.z)), fmaxf(w1z, w2z)));
        perm[k] = k; follow[k] = 0;
    }
    // explicit stack of ranges; where = index of the int that receives the pointer to the subtree (node * 4 + child) in nodes_new rows [3]
    int st_a[SPLIT_MAXT], st_b[SPLIT_MAXT], st_where[SPLIT_MAXT], sp = 0;
    int next_node = n_inner + offsets[i];
    st_a[0] = 0; st_b[0] = cnt; st_where[0] = 4 * (i >> 1) + (i & 1); sp = 1;
    while (sp > 0) {
        --sp;
        const int a = st_a[sp], b = st_b[sp], where = st_where[sp], n = b - a;
        float3 plo = lo[perm[a]], phi = hi[perm[a]];
        for (int k = a + 1; k < b; ++k) {
            plo = make_float3(fminf(plo.x, lo[perm[k]].x), fminf(plo.y, lo[perm[k]].y), fminf(plo.z, lo[perm[k]].z));
            phi = make_float3(fmaxf(phi.x, hi[perm[k]].x), fmaxf(phi.y, hi[perm[k]].y), fmaxf(phi.z, hi[perm[k]].z));
        }
        const float parent_area = box_half_area(plo, phi);
        // best object split: the references sorted by box centre along each axis, every cut position (SAH, costs 1 : 1)
        float best = 3.0e38f; int best_axis = -1, best_k = 0;
        if (n > 1) {
            for (int axis = 0; axis < 3; ++axis) {
                for (int x = a + 1; x < b; ++x) {                                   // insertion sort of perm[a..b) by centre
                    const int p = perm[x];
                    const float key = axis == 0 ? lo[p].x + hi[p].x : (axis == 1 ? lo[p].y + hi[p].y : lo[p].z + hi[p].z);
                    int y = x - 1;
                    while (y >= a) {
                        const int q = perm[y];
                        const float kq = axis == 0 ? lo[q].x + hi[q].x : (axis == 1 ? lo[q].y + hi[q].y : lo[q].z + hi[q].z);
                        if (kq <= key) break;
                        perm[y + 1] = q; --y;
                    }
                    perm[y + 1] = p;
                }
                float3 rl = lo[perm[b - 1]], rh = hi[perm[b - 1]];
                for (int k = b - 1; k > a; --k) {
                    rl = make_float3(fminf(rl.x, lo[perm[k]].x), fminf(rl.y, lo[perm[k]].y), fminf(rl.z, lo[perm[k]].z));
                    rh = make_float3(fmaxf(rh.x, hi[perm[k]].x), fmaxf(rh.y, hi[perm[k]].y), fmaxf(rh.z, hi[perm[k]].z));
                    right_area[k] = box_half_area(rl, rh);
                }
                float3 ll = lo[perm[a]], lh = hi[perm[a]];
                for (int k = a + 1; k < b; ++k) {                                   // left = [a, k), right = [k, b)
                    const float cost = box_half_area(ll, lh) * (float)(k - a) + right_area[k] * (float)(b - k);
                    if (cost < best) { best = cost; best_axis = axis; best_k = k; }
                    ll = make_float3(fminf(ll.x, lo[perm[k]].x), fminf(ll.y, lo[perm[k]].y), fminf(ll.z, lo[perm[k]].z));
                    lh = make_float3(fmaxf(lh.x, hi[perm[k]].x), fmaxf(lh.y, hi[perm[k]].y), fmaxf(lh.z, hi[perm[k]].z));
                }
            }
        }
        if (n > 1 && best_axis < 0) { best_axis = 0; best_k = a + n / 2; }          // no finite cost (non-finite vertices): any valid cut keeps the tree well-formed
        // SplitBVHBuilder.cpp:142: a leaf where the leaf's SAH is the minimum AND it is small enough
        const bool leaf = n == 1 || (n <= max_leaf && !(parent_area + best < parent_area * (float)n));
        if (leaf) {
            reinterpret_cast<int*>(nodes_new + 4 * (size_t)(where >> 2) + 3)[where & 3] = ~(base + a);
            for (int k = a; k < b; ++k) follow[k] = b - 1 - k;
            continue;
        }
        for (int x = a + 1; x < b; ++x) {                                           // back to the winning axis' order
            const int p = perm[x];
            const float key = best_axis == 0 ? lo[p].x + hi[p].x : (best_axis == 1 ? lo[p].y + hi[p].y : lo[p].z + hi[p].z);
            int y = x - 1;
            while (y >= a) {
                const int q = perm[y];
                const float kq = best_axis == 0 ? lo[q].x + hi[q].x : (best_axis == 1 ? lo[q].y + hi[q].y : lo[q].z + hi[q].z);
                if (kq <= key) break;
                perm[y + 1] = q; --y;
            }
            perm[y + 1] = p;
        }
        float3 l0 = lo[perm[a]], h0 = hi[perm[a]], l1 = lo[perm[best_k]], h1 = hi[perm[best_k]];
        for (int k = a + 1; k < best_k; ++k) {
            l0 = make_float3(fminf(l0.x, lo[perm[k]].x), fminf(l0.y, lo[perm[k]].y), fminf(l0.z, lo[perm[k]].z));
            h0 = make_float3(fmaxf(h0.x, hi[perm[k]].x), fmaxf(h0.y, hi[perm[k]].y), fmaxf(h0.z, hi[perm[k]].z));
        }
        for (int k = best_k + 1; k < b; ++k) {
            l1 = make_float3(fminf(l1.x, lo[perm[k]].x), fminf(l1.y, lo[perm[k]].y), fminf(l1.z, lo[perm[k]].z));
            h1 = make_float3(fmaxf(h1.x, hi[perm[k]].x), fmaxf(h1.y, hi[perm[k]].y), fmaxf(h1.z, hi[perm[k]].z));
        }
        const int id = next_node++;
        nodes_new[4 * (size_t)id + 0] = make_float4(l0.x, h0.x, l0.y, h0.y);       // the packed node layout, gpt_scene.cuh
        nodes_new[4 * (size_t)id + 1] = make_float4(l1.x, h1.x, l1.y, h1.y);
        nodes_new[4 * (size_t)id + 2] = make_float4(l0.z, h0.z, l1.z, h1.z);
        nodes_new[4 * (size_t)id + 3] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        reinterpret_cast<int*>(nodes_new + 4 * (size_t)(where >> 2) + 3)[where & 3] = 4 * id;
        st_a[sp] = best_k; st_b[sp] = b; st_where[sp] = 4 * id + 1; ++sp;
        st_a[sp] = a; st_b[sp] = best_k; st_where[sp] = 4 * id + 0; ++sp;
    }
    for (int k = 0; k < cnt; ++k) {                                                 // the leaf's rows in their new order
        const size_t src = GPT_TRI_STRIDE * (size_t)(base + perm[k]), dst = GPT_TRI_STRIDE * (size_t)(base + k);
        float4 c = tris_old[src + GPT_TRI_B + 1];
        c.w = __int_as_float(follow[k]);
        tris_new[dst] = tris_old[src]; tris_new[dst + GPT_TRI_B] = tris_old[src + GPT_TRI_B]; tris_new[dst + GPT_TRI_B + 1] = c;
    }
}

// status: 0 ok (out filled: a NEW node array and a NEW triangle array of the same triangle count), 1 nothing to split, 2 too deep
cudaError_t gptk_split_leaves(const float4* nodes, int n_inner, const float4* tris, int n_tris, const float4* ref_tris, int max_leaf,
                              RepackResult* out, int* status, unsigned long long* launches, cudaStream_t st)
{
    *out = RepackResult{}; *status = 1;
    if (n_inner < 1 || n_tris < 1) return cudaSuccess;
    if (max_leaf < 1) max_leaf = 1;
    const int n_child = 2 * n_inner;
    int *counts = nullptr, *offsets = nullptr, *flag = nullptr, *level = nullptr;
    void* tmp = nullptr; size_t tmp_bytes = 0;
    int extra = 0, depth = 1;
    cudaError_t e;
#define CK(x) do { e = (x); if (e != cudaSuccess) goto fail; } while (0)
    CK(cudaMalloc(&counts, sizeof(int) * (size_t)(n_child + 1)));
    CK(cudaMalloc(&offsets, sizeof(int) * (size_t)(n_child + 1)));
    CK(cudaMalloc(&flag, sizeof(int)));
    CK(cudaMemsetAsync(counts, 0, sizeof(int) * (size_t)(n_child + 1), st));
    k_split_count<<<(n_child + 255) / 256, 256, 0, st>>>(nodes, n_inner, tris, max_leaf, counts);
    CK(cudaGetLastError());
    CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, counts, offsets, n_child + 1, st));
    CK(cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 16));
    CK(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, counts, offsets, n_child + 1, st));
    CK(cudaMemcpyAsync(&extra, offsets + n_child, sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    *launches += 3;
    if (extra <= 0) goto done;
    {
        const int n_new = n_inner + extra;                                      // an upper bound: unused rows stay zero and unreachable
        CK(cudaMalloc(&out->nodes, sizeof(float4) * 4 * (size_t)n_new));
        CK(cudaMalloc(&out->tris, sizeof(float4) * GPT_TRI_STRIDE * ((size_t)n_tris + 1)));
        CK(cudaMemsetAsync(out->nodes, 0, sizeof(float4) * 4 * (size_t)n_new, st));
        CK(cudaMemcpyAsync(out->nodes, nodes, sizeof(float4) * 4 * (size_t)n_inner, cudaMemcpyDeviceToDevice, st));
        CK(cudaMemcpyAsync(out->tris, tris, sizeof(float4) * GPT_TRI_STRIDE * ((size_t)n_tris + 1), cudaMemcpyDeviceToDevice, st));
        k_split_build<<<(n_child + 63) / 64, 64, 0, st>>>(nodes, n_inner, tris, ref_tris, offsets, max_leaf, out->nodes, out->tris);
        CK(cudaGetLastError());
        *launches += 1;
        // depth of the new tree (children sit behind their parents here too)
        CK(cudaMalloc(&level, sizeof(int) * (size_t)n_new));
        CK(cudaMemsetAsync(level, 0, sizeof(int) * (size_t)n_new, st));
        { const int one = 1; CK(cudaMemcpyAsync(level, &one, sizeof(int), cudaMemcpyHostToDevice, st)); }
        for (; depth <= GPT_STACK_SIZE; ++depth) {
            int h_changed = 0;
            CK(cudaMemsetAsync(flag, 0, sizeof(int), st));
            k_node_levels<<<(n_new + 255) / 256, 256, 0, st>>>(out->nodes, n_new, level, depth, flag);
            CK(cudaGetLastError());
            CK(cudaMemcpyAsync(&h_changed, flag, sizeof(int), cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            if (!h_changed) break;
        }
        out->n_nodes = n_new; out->n_tris = n_tris; out->depth = depth;
        *status = depth >= GPT_STACK_SIZE ? 2 : 0;
        if (*status) { cudaFree(out->nodes); cudaFree(out->tris); *out = RepackResult{}; }
    }
done:
    cudaFree(counts); cudaFree(offsets); cudaFree(flag); cudaFree(tmp); cudaFree(level);
    return cudaSuccess;
fail:
    cudaFree(counts); cudaFree(offsets); cudaFree(flag); cudaFree(tmp); cudaFree(level);
    cudaFree(out->nodes); cudaFree(out->tris); *out = RepackResult{};
    return e;
#undef CK
}
