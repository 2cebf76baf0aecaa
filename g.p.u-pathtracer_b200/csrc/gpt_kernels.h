// gpt_kernels.h -- host-callable launchers of the device code (internal; the public surface is include/gpt.h)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "gpt_scene.cuh"
#include <vector>

// Optional per-kernel-class device timing (GPT_FLAG_TIME_KERNELS): one CUDA event is recorded on the launching
// stream in front of every launch (and one after the last); the interval that follows an event is attributed to
// the class of the kernel launched right after it.  Read back with gpt_get_kernel_times after a sync.
enum { KT_GENERATE = 0, KT_EXTEND = 1, KT_SHADE = 2, KT_COLLECT = 3, KT_ACCUMULATE = 4, KT_RESOLVE = 5, KT_OTHER = 6, KT_END = 7, KT_CLASSES = 8 };
struct KernelTimer {
    std::vector<cudaEvent_t> ev;
    std::vector<int> kind;
    size_t used = 0;
    void tick(int k, cudaStream_t st)
    {
        if (used == ev.size()) {
            cudaEvent_t e;
            if (cudaEventCreate(&e) != cudaSuccess) return;
            ev.push_back(e); kind.push_back(KT_END);
        }
        kind[used] = k;
        cudaEventRecord(ev[used], st);
        ++used;
    }
    // sums the finished intervals into ms[]/n[] and forgets them; the caller has synchronised
    void collect(double* ms, unsigned long long* n)
    {
        for (size_t i = 0; i + 1 < used; ++i) {
            if (kind[i] == KT_END) continue;
            float t = 0.0f;
            if (cudaEventElapsedTime(&t, ev[i], ev[i + 1]) == cudaSuccess) { ms[kind[i]] += t; n[kind[i]] += 1; }
        }
        used = 0;
    }
    void release() { for (cudaEvent_t e : ev) cudaEventDestroy(e); ev.clear(); kind.clear(); used = 0; }
};

cudaError_t gptk_render_mega(const DevScene& sc, const RenderParams& p, float4* accum_f4, unsigned long long sample_begin,
                             uint32_t sample_count, unsigned int* work_counter, unsigned long long* stats,
                             int sm_count, cudaStream_t st);
cudaError_t gptk_render_legacy(const DevScene& sc, const RenderParams& p, float* accum_v3, uint32_t* draw,
                               unsigned long long constant_pdf, unsigned int* work_counter, unsigned long long* stats,
                               int sm_count, cudaStream_t st);
cudaError_t gptk_primary(const DevScene& sc, const RenderParams& p, unsigned long long sample,
                         int* geom, int* tri_idx, int* prim_idx, float* t, int sm_count, cudaStream_t st);
cudaError_t gptk_trace_rays(const DevScene& sc, const float* o, const float* d, size_t n, int cull,
                            int* tri_idx, int* tri_addr, float* t, int sm_count, cudaStream_t st);
cudaError_t gptk_resolve(const float4* accum, int n, uint32_t* rgba8, float* rgb, int sm_count, cudaStream_t st);

// root-side fused peer reduce + running-sum update + resolve; frames[] = device-accessible pointers (own or IPC-mapped)
cudaError_t gptk_reduce_resolve(const float4* const* frames, int n_frames, int n_pix, float4* total, uint32_t* rgba8, float* rgb,
                                int sm_count, cudaStream_t st);

// device re-pack (gpt_repack.cu): reference arrays (device pointers) -> packed nodes / triangles.
// *out_nodes / *out_tris are cudaMalloc'd by the callee; returns the packed triangle count.
struct RepackResult { float4* nodes; float4* tris; int n_nodes; int n_tris; int bad_layout; int depth; };   // bad_layout 5 = deeper than the traversal stack
cudaError_t gptk_repack(const float4* ref_nodes, size_t n_node_vec4, const float4* ref_tris, size_t n_tri_vec4,
                        const int* ref_indices, size_t n_indices, RepackResult* out, cudaStream_t st);

// A second packed tree over the same triangle references, made by the device builder below from the caller's exact vertices
// (gpt_repack.cu).  packed_tris / n_tris = the first packing; ref_tris = the caller's triangle array it was packed from.
// *status: 0 ok, 1 builder refused / too deep, 2 internal mismatch, 3 fewer than 2 triangles.  Synchronises `st`.
cudaError_t gptk_rebuild(const float4* packed_tris, int n_tris, const float4* ref_tris, int max_leaf, int ploc_radius,
                         RepackResult* out, int* status, float* build_ms, unsigned long long* launches, cudaStream_t st);

// The packed tree with every leaf of max_leaf + 1 .. 16 references replaced by a small SAH subtree (gpt_repack.cu): new node and
// triangle arrays, same triangle count, rows only permuted inside their leaf.  *status: 0 ok, 1 nothing to split, 2 too deep.
cudaError_t gptk_split_leaves(const float4* nodes, int n_inner, const float4* tris, int n_tris, const float4* ref_tris, int max_leaf,
                              RepackResult* out, int* status, unsigned long long* launches, cudaStream_t st);

// device-side LBVH builder (gpt_build.cu): indexed mesh (device pointers) -> arrays in CudaBVH's compact layout,
// cudaMalloc'd by the callee.  status: 0 ok, 1 vertex index out of range, 2 non-finite vertex, 3 too large, 4 too deep
struct BuildResult { float4* nodes; size_t n_node_vec4; float4* tris; size_t n_tri_vec4; int* indices; size_t n_indices;
                     int n_leaves; int depth; int status; int launches; };
// ploc_radius: 0 = Morton radix tree (Karras 2012); > 0 = PLOC clustering with that search radius (<= 32).
// status: 1 index out of range, 2 non-finite vertex, 3 too large, 4 too deep, 5 clustering did not converge
cudaError_t gptk_build_lbvh(const float* verts, size_t n_verts, const int* tri, size_t n_tris, int max_leaf, int ploc_radius,
                            BuildResult* out, cudaStream_t st);

// device peaks measured by micro-benchmarks (gpt_peaks.cu; diagnostics for bench.py's roofline)
cudaError_t gptk_measure_peaks(int sm_count, double* l1_gbs, double* l2_gbs, double* hbm_gbs, double* warp_inst_per_s,
                               double* stream_l1_gbs, double* stream_l2_gbs);

// 4-wide quantised BVH collapsed on the device from the packed binary nodes (gpt_wide.cu).
// status: 0 ok, 1 non-finite / inverted box (not quantisable), 2 internal, 3 not a tree
struct WideResult { float4* wnodes; int n_wide; int levels; int status; int launches; };
cudaError_t gptk_build_wide(const float4* nodes2, int n_inner, WideResult* out, cudaStream_t st);

// wavefront pipeline (gpt_wavefront.cu).  Per-path arrays hold `capacity` entries, lsum one per pixel.
struct WaveBuffers {
    // Records of the paths alive at a bounce lie in QUEUE order, double-buffered by bounce parity (gpt_wavefront.cu):
    // ray: 2 x (2 float4 per slot) = origin + direction, whose w lanes carry the colour mask's x / y; mask_z: 2 x (1 float per slot);
    // queue[2]: the slot's path id; hit: 1 int2 per slot (written by the extend kernel, read by the shade kernel of the same bounce);
    // accu: by path id
    float4 *ray, *accu, *lsum;
    float* mask_z;
    int2* hit;
    uint32_t* queue[2];
    unsigned int* counters;     // gptk_wave_counter_count(depth) entries
    size_t capacity;            // paths the per-path arrays hold (lsum: one per pixel of the frame)
    // optional ray reordering before each secondary extend (GPT_WF_SORT): all nullptr = off
    uint32_t* sort_keys[2]; uint32_t* sort_vals; void* sort_tmp; size_t sort_tmp_bytes;
};
size_t gptk_wave_counter_count(unsigned int depth);
size_t gptk_wave_sort_tmp_bytes(size_t capacity);
// per-context launch configuration of the pipeline (grids from the kernels' occupancy on this device, environment knobs)
struct WaveLaunchCfg { int ready, sm_count, ext_grid, wide_grid, shade_grid, sort_mode, refill_min, node_k_env; size_t wave_cap_env; };
void gptk_wave_launch_cfg(WaveLaunchCfg* cfg, int sm_count);
// wb.counters must be all zero on entry and are all zero again once the enqueued work has run
cudaError_t gptk_render_wavefront(const DevScene& sc, const RenderParams& p, const WaveBuffers& wb, const WaveLaunchCfg& cfg,
                                  float4* accum_f4, float* accum_v3, uint32_t* draw, unsigned long long constant_pdf,
                                  unsigned long long sample_begin, uint32_t sample_count,
                                  unsigned long long* stats, cudaStream_t st, unsigned long long* launches,
                                  KernelTimer* kt);
