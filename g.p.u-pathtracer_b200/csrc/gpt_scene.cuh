// gpt_scene.cuh -- device-side scene and launch-parameter structs.
//
// HBM layout after the device re-pack (gpt_repack.cu), replacing the three arrays the reference
// uploads verbatim (BasicScene.cpp:294-313; layouts CudaBVH.cpp:196-224):
//   nodes : float4[4] per inner node (64 B, 64-B aligned)
//           [0] = c0.lo.x c0.hi.x c0.lo.y c0.hi.y     (unchanged from CudaBVH.cpp:219)
//           [1] = c1.lo.x c1.hi.x c1.lo.y c1.hi.y     (:220)
//           [2] = c0.lo.z c0.hi.z c1.lo.z c1.hi.z     (:221)
//           [3] = int child0, child1, 0, 0            child >= 0: float4 index of an inner node
//                                                     child <  0: ~(index of the leaf's first packed triangle)
//   tris  : float4[3] per triangle reference (48 B), leaf after leaf, NO terminator slots
//           [0] = v0.xyz , bits(original OBJ face index)      (the remap gpuTriIndices[hit], cudaUtils.h:452-456)
//           [1] = (v1-v0).xyz , bits(slot address in the caller's array) (the reference's pre-remap hitIndex)
//           [2] = (v2-v0).xyz , bits(number of triangles that FOLLOW in the same leaf; 0 = last)  (replaces the 0x80000000
//                 terminator read, :413; the first triangle of a leaf therefore carries leaf size - 1)
//   spheres / boxes : the caller's 44-B / 52-B structs copied verbatim, staged to shared memory per CTA.
#pragma once
#include "gpt_math.cuh"

#define GPT_SENTINEL 0x76543210          // bottom-of-stack marker (same value as cudaUtils.h:21)
#define GPT_STACK_SIZE 64                // cudaUtils.h:22; builder MaxDepth 64 (SplitBVHBuilder.hpp:15)
#define GPT_W4_EMPTY 0x7fffffff           // unused child slot of a wide node (gpt_wide.cu)
#define GPT_W4_STACK 96                  // traversal stack of the wide tree: <= 3 pushes per wide level
#define GPT_MAX_SMEM_SPHERES 64
#define GPT_MAX_SMEM_BOXES 32

enum { GEO_TRI = 0, GEO_SPHERE = 1, GEO_BOX = 2, GEO_NONE = 3 };
enum { MAT_DIFF = 0, MAT_METAL = 1, MAT_SPEC = 2, MAT_REFR = 3 };
enum { RP_NO_JITTER = 1u, RP_METAL_FIXED = 2u, RP_COUNT_BOUNCES = 8u, RP_FAST = 16u };

struct DevScene {
    const float4* nodes;     // packed nodes (see above); nullptr/0 => no mesh
    const float4* tris;      // packed triangles
    const float*  spheres;   // n_spheres * 11 floats
    const float*  boxes;     // n_boxes * 13 floats
    int n_nodes;             // inner nodes
    int n_tris;
    int n_spheres;
    int n_boxes;
    // N4: per-face materials (nullptr / 0 = the reference's single global triangle material, tracer.cu:131-135)
    const float* tri_mats;   // n_tri_mats * 7 floats: col rgb, emi rgb, bits(Mat)
    const int*   face_mat;   // n_faces entries, indexed by the remapped OBJ face index (cudaUtils.h:452-456)
    int n_tri_mats;
    int n_faces;
    // N3: optional 4-wide quantised tree over the same packed triangles (gpt_wide.cu); nullptr = traverse the binary tree
    const float4* wnodes;
};

// the kernelInfo fields the device code reads (CpuStructs.hpp:45-72), flattened
struct RenderParams {
    int   width, height;
    float front[3], right[3], up[3], pos[3];
    float dist, aspect, fov;
    int   cull;
    unsigned int depth;
    float air, glass;
    int   tri_mat;
    float bk[3];
    float phong;
    float emi[3], col[3];
    unsigned long long seed;
    unsigned int flags;
};

__device__ __forceinline__ f3 ld3(const float* p) { return mk3(p[0], p[1], p[2]); }

// Packed-triangle addressing: 48 B per triangle, rows (v0 | e1 | e2) = three LDG.128; the two edge rows first -- the determinant
// test (cudaUtils.h:145-153) needs only them.  (A 64-byte layout whose edge rows come with one 256-bit load was measured 1-2 %
// slower on every scene: profiles/r02_schedules.md #52.)
#define GPT_TRI_STRIDE 3
#define GPT_TRI_B 1
__device__ __forceinline__ float4 tri_row_a(const float4* __restrict__ tris, const int ti) { return __ldg(tris + GPT_TRI_STRIDE * (size_t)ti); }
__device__ __forceinline__ void tri_rows_bc(const float4* __restrict__ tris, const int ti, float4& b, float4& c)
{
    b = __ldg(tris + GPT_TRI_STRIDE * (size_t)ti + 1); c = __ldg(tris + GPT_TRI_STRIDE * (size_t)ti + 2);
}
