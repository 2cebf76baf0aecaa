// TEST INFRASTRUCTURE ONLY (oracle/): thin extern "C" wrapper that drives the
// REFERENCE's own host BVH builder, compiled in place from /root/reference (never copied):
//   GpuPathTracer/{BVH,BVHNode,bvh_util,SplitBVHBuilder,Sort,CudaBVH}.cpp
// following the call sequence of GpuPathTracer/BasicScene.cpp:278-318
//   SceneMesh scene(nT,nV,tris,verts); BVH bvh(&scene, Platform(), BuildParams());
//   CudaBVH gpu(bvh, BVHLayout_Compact); getGpuNodes()/getDebugTri()/getGpuTriWoop()/getGpuTriIndices()
// Output: oracle/_ref/librefbvh.so (git-ignored). Product code never links this.
#include "BVH.hpp"
#include "CudaBVH.hpp"
#include "SceneMesh.hpp"
#include <cstdlib>
#include <cstring>
#include <vector>

extern "C" {

struct refbvh_result {
    void* nodes;      unsigned int n_node_vec4;   // CudaBVH.hpp:58,63
    void* tris;       unsigned int n_tri_vec4;    // getDebugTri(), CudaBVH.hpp:60,65
    void* woop;       unsigned int n_woop_vec4;   // getGpuTriWoop(), CudaBVH.hpp:59,64
    int*  indices;    unsigned int n_indices;     // getGpuTriIndices(), CudaBVH.hpp:61,66
    unsigned int n_leaves;
};

// verts: nv*3 floats, tri_idx: nt*3 ints. Returns 0 on success.
int refbvh_build(const float* verts, int nv, const int* tri_idx, int nt, refbvh_result* out)
{
    if (!verts || !tri_idx || !out || nt < 2 || nv < 3) return -1; // 1-tri root-leaf asserts (CudaBVH.cpp:141)
    std::vector<glm::vec3> ve((size_t)nv);
    for (int i = 0; i < nv; ++i) ve[i] = glm::vec3(verts[3*i], verts[3*i+1], verts[3*i+2]);
    std::vector<SceneMesh::Triangle> tr;
    tr.reserve((size_t)nt);
    for (int i = 0; i < nt; ++i)
        tr.push_back(SceneMesh::Triangle(glm::ivec3(tri_idx[3*i], tri_idx[3*i+1], tri_idx[3*i+2])));
    SceneMesh scene(nt, nv, tr, ve);
    Platform platform;
    BVH::BuildParams params;
    params.enablePrints = false;
    BVH bvh(&scene, platform, params);
    CudaBVH* g = new CudaBVH(bvh, BVHLayout_Compact);
    out->nodes = g->getGpuNodes();        out->n_node_vec4 = g->getGpuNodesSize();
    out->tris = g->getDebugTri();         out->n_tri_vec4 = g->getDebugTriSize();
    out->woop = g->getGpuTriWoop();       out->n_woop_vec4 = g->getGpuTriWoopSize();
    out->indices = g->getGpuTriIndices(); out->n_indices = g->getGpuTriIndicesSize();
    out->n_leaves = g->getLeafnodeCount();
    delete g; // dtor is empty (CudaBVH.cpp:68-70); malloc'd arrays survive and are freed below
    return 0;
}

void refbvh_free(refbvh_result* r)
{
    if (!r) return;
    free(r->nodes); free(r->tris); free(r->woop); free(r->indices);
    memset(r, 0, sizeof(*r));
}

} // extern "C"
