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
