/*
 * include/gpt.h -- C-ABI of the B200-native progressive path-tracing hot path.
 *
 * Drop-in boundary for G.P.U-PathTracer's render launch:
 *   reference call site   BasicScene::run -> launchKernel(info)     GpuPathTracer/BasicScene.cpp:395-404
 *   reference definition  void BasicScene::launchKernel(const kernelInfo&)  GpuPathTracer/tracer.cu:405-415
 *   reference data        kernelInfo / CamInfo / BVHData            GpuPathTracer/CpuStructs.hpp:19-72
 *                         Sphere / Box / GeoType / Mat              GpuPathTracer/CommomStructs.hpp:11-82
 *                         node / triangle / index arrays            GpuPathTracer/CudaBVH.hpp:58-66, CudaBVH.cpp:121-270
 *
 * Plain C: pointers and sizes only, no C++/torch types.  Every entry returns a gpt_status
 * (0 = ok, negative = error) and never calls exit() (the reference's checkCudaErrors does,
 * utilfun.hpp:81-90).
 * There is NO CPU fallback: without a CUDA device every compute entry returns GPT_ERR_CUDA.
 *
 * Streams and threads.  A context owns ONE set of scratch buffers (path state, ray queues and their counters, work
 * cursor, launch sums): ONE stream and ONE host thread may have work of a context in flight at a time, like the
 * reference's single frame in flight (BasicScene.cpp:395).  Use one context per stream / thread for concurrency;
 * the library is re-entrant across contexts.  gpt_launch / gpt_render / gpt_resolve / gpt_reduce_resolve only enqueue work
 * on `stream` EXCEPT that a call may block the host (cudaDeviceSynchronize / cudaStreamSynchronize) when it
 *   - has to grow the scratch buffers (first call, a larger frame, more samples per call, a deeper path),
 *   - finds new BVH pointers in kernelInfo and re-packs them (gpt_launch; this also builds the rebuilt tree, gpt_rebuild_info), or
 *     builds the 4-wide tree (first render of a scene),
 *   - is one of the timed calls of the per-scene trial (gpt_trial_info; the first <= 15 render calls of >= 1 Mi paths --
 *     >= 393 216 paths when a second tree is a candidate -- after a
 *     scene upload on scenes with more than one candidate tree or ray order; likewise the first <= 9 small calls of >= 16 Ki paths
 *     for the megakernel's own tree trial, gpt_megakernel_tree_info; GPT_FLAG_NO_TRIAL switches both off).
 * gpt_upload_scene, gpt_build_bvh, gpt_set_triangle_materials, gpt_get_stats, gpt_get_kernel_times, gpt_reset_stats,
 * gpt_render_frame_host and the gpt_frame_* setup calls always synchronise.
 */
#ifndef GPT_H
#define GPT_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define GPT_API
#else
#define GPT_API __attribute__((visibility("default")))
#endif

typedef enum {
    GPT_OK = 0,
    GPT_ERR_INVALID = -1,     /* bad argument */
    GPT_ERR_CUDA = -2,        /* CUDA runtime error; see gpt_last_error_string */
    GPT_ERR_NO_SCENE = -3,    /* launch before gpt_upload_scene / without BVH pointers */
    GPT_ERR_LAYOUT = -4,      /* malformed node/triangle arrays */
    GPT_ERR_IO = -5
} gpt_status;

/* CommomStructs.hpp:11-12 -- the enum ORDER is ABI (Mat::METAL == 1, SPEC == 2) */
typedef enum { GPT_GEO_TRI = 0, GPT_GEO_SPHERE = 1, GPT_GEO_BOX = 2, GPT_GEO_NONE = 3 } gpt_geo_type;
typedef enum { GPT_MAT_DIFF = 0, GPT_MAT_METAL = 1, GPT_MAT_SPEC = 2, GPT_MAT_REFR = 3 } gpt_mat;

/* Sphere, CommomStructs.hpp:18-39 (44 B) */
typedef struct { float posRad[4]; float emi[3]; float col[3]; int32_t mat; } gpt_sphere;
/* Box, CommomStructs.hpp:42-48 (52 B).  A stub in the reference (intersect returns 0);
 * implemented here from its commented spec cudaUtils.h:31-77. */
typedef struct { float mn[3]; float mx[3]; float emi[3]; float col[3]; int32_t mat; } gpt_box;

/* CamInfo, CpuStructs.hpp:19-29 (84 B) */
typedef struct {
    float front[3], right[3], up[3], pos[3];
    float dist;
    float pitch, yaw;
    float aspect;
    float fov;
    uint8_t dirty; uint8_t _pad0[3];
    float bias[2];
    uint8_t enabled; uint8_t _pad1[3];
} gpt_cam_info;

/* BVHData, CpuStructs.hpp:31-42 (48 B): device pointers + element counts (vec4s / ints) */
typedef struct {
    void*   dev_triNode;       size_t triNodeSize;
    int32_t* dev_triIndicesPtr; size_t triIndicesSize;
    void*   dev_triPtr;        size_t triSize;
} gpt_bvh_data;

/* kernelInfo, CpuStructs.hpp:45-72 (296 B) -- bit-compatible, so a maintainer's adapter is
 *   void BasicScene::launchKernel(const kernelInfo& i){ gpt_launch(ctx,(const gpt_kernel_info*)&i,0); } */
typedef struct {
    uint32_t blockSize[3]; uint32_t _pad0;   /* dim3; accepted and ignored (persistent grid) */
    uint64_t hash;                           /* per-frame seed, uf::hash(frame) utilfun.cpp:380-389 */
    uint32_t* dev_drawRes;                   /* W*H RGBA8 0x00BBGGRR, row 0 = bottom */
    float*   accumBuffer;                    /* W*H vec3 running mean (legacy contract) */
    int32_t  width, height;
    gpt_cam_info cam;
    uint32_t _pad1;
    void*    triangleTex;  size_t numVerts;  /* dead brute-force input; ignored */
    gpt_sphere* sphereTex; size_t numSpheres;
    uint8_t  cullBackFaces; uint8_t _pad2[3];
    uint32_t depth;
    uint64_t constantPdf;                    /* N = samples since last reset */
    gpt_bvh_data bvhData;
    float    time_elapsed;                   /* ignored (never written in the reference either) */
    float    air_ref_index, glass_ref_index;
    int32_t  triCurrentMat;                  /* gpt_mat */
    uint8_t  clearBuffer; uint8_t _pad3[3];
    float    bkColor[3];
    float    phongExpo;
    float    emi[3], col[3];
    uint32_t _pad4;
} gpt_kernel_info;

/* context flags (gpt_set_flags) */
enum {
    GPT_FLAG_NO_JITTER   = 1u << 0,  /* camera jitter off (hit-ID parity runs) */
    GPT_FLAG_METAL_FIXED = 1u << 1,  /* use the reflection vector w1 instead of the reference's int `w` */
    GPT_FLAG_MEGAKERNEL  = 1u << 2,  /* always the single persistent megakernel (default: only for calls of fewer than
                                      * GPT_SMALL_FRAME_PATHS = 393216 paths, where launch overhead dominates) */
    GPT_FLAG_COUNT_BOUNCES = 1u << 3, /* fill gpt_stats.rays_by_bounce (slower) */
    GPT_FLAG_TIME_KERNELS = 1u << 4, /* CUDA-event timing per kernel class, read with gpt_get_kernel_times */
    /* Traversal structure of the wavefront pipeline and of gpt_trace_rays.  Default: the binary tree as
     * uploaded (CudaBVH.cpp:121-270; visit order and therefore ties identical to cudaUtils.h:298-448).  Alternative: a
     * 4-wide tree with 8-bit quantised child boxes collapsed from it on the device (64 B per four children).  Same
     * nearest hit (t, face index); between two copies of a spatially split triangle -- bit-equal t -- the reported slot
     * may be the other copy.  Measured on B200: slower than the binary tree on small scenes and on compact
     * objects in a box, faster on scenes that fill space and on anything large.  Unless one of the two flags or
     * GPT_WIDE_MIN_TRIS in the environment decides, scenes from 64 Ki packed triangles up are put on a timed trial during
     * their first large render calls (gpt_trial_info); until it ends the wide tree is used from 1 Mi triangles. */
    GPT_FLAG_WIDE_BVH   = 1u << 5,   /* always the 4-wide tree */
    GPT_FLAG_BINARY_BVH = 1u << 6,   /* always the binary tree */
    GPT_FLAG_WAVEFRONT  = 1u << 7,   /* always the wavefront pipeline, also for small calls */
    GPT_FLAG_BUILD_LBVH = 1u << 8,   /* gpt_build_bvh: plain Morton radix tree instead of the (default) PLOC clustering */
    GPT_FLAG_FAST       = 1u << 10,  /* OPT-IN fast shading arithmetic, held to north_star's statistical tolerance (1024-spp image against
                                      * the reference's GPU kernel: per-pixel z-scores, box-filtered RMSE <= 1e-3) instead of bit-exactness
                                      * with the oracle: rsqrt normalisation, hardware sin/cos/pow/reciprocal (MUFU) in the shade kernel.
                                      * Nearest hits -- t, face index, the primary-hit map -- are unchanged (the traversal is not touched).
                                      * Off by default: gpt_launch / gpt_render stay bit-exact unless the caller sets it. */
    GPT_FLAG_NO_TRIAL   = 1u << 9,   /* never run the timed per-scene trial (gpt_trial_info): no render call then synchronises, so
                                      * calls are strictly asynchronous and can be captured into a CUDA graph; the size defaults
                                      * decide tree and ray order (also: GPT_NO_TRIAL=1 in the environment) */
    GPT_FLAG_KEEP_TREE  = 1u << 11   /* traverse the BVH exactly as uploaded (or its 4-wide collapse): do not build a second tree over
                                      * the same triangles with this library's builder (gpt_rebuild_info; also GPT_REBUILD=0 in the
                                      * environment).  Must be set before the upload / the first gpt_launch to save the build. */
};

typedef struct {
    uint64_t rays;                /* bounce-loop iterations executed since the last gpt_reset_stats */
    uint64_t rays_by_bounce[16];  /* only with GPT_FLAG_COUNT_BOUNCES */
    uint64_t paths;               /* (pixel,sample) paths started */
    uint64_t kernel_launches;     /* library kernels launched */
} gpt_stats;

/* Device time per kernel class since the last gpt_reset_stats (GPT_FLAG_TIME_KERNELS): CUDA events recorded on the
 * launching stream around every launch (replaces the reference's commented-out uf::GpuTimer, BasicScene.cpp:402-418). */
enum { GPT_K_GENERATE = 0, GPT_K_EXTEND = 1, GPT_K_SHADE = 2, GPT_K_COLLECT = 3, GPT_K_ACCUMULATE = 4, GPT_K_RESOLVE = 5, GPT_K_OTHER = 6 };
typedef struct { double ms[8]; uint64_t launches[8]; } gpt_kernel_times;

typedef struct gpt_ctx gpt_ctx;
typedef void* gpt_stream;   /* cudaStream_t */

/* ---- lifetime ---- */
GPT_API int  gpt_create(gpt_ctx** out, int device);          /* replaces uf::findCudaDevice, utilfun.cpp:72-99 */
GPT_API void gpt_destroy(gpt_ctx* ctx);
GPT_API int  gpt_set_flags(gpt_ctx* ctx, uint32_t flags);
GPT_API const char* gpt_last_error_string(gpt_ctx* ctx);     /* ctx may be NULL: last create error */
GPT_API const char* gpt_version(void);
GPT_API int  gpt_sync(gpt_ctx* ctx, gpt_stream stream);      /* cudaStreamSynchronize, BasicScene.cpp:395 */

/* ---- scene ---- */
/* Takes the arrays CudaBVH returns (getGpuNodes/getDebugTri/getGpuTriIndices, CudaBVH.hpp:58-66),
 * as host (on_device=0) or device (on_device=1) pointers; replaces the cudaMalloc+cudaMemcpy block
 * BasicScene.cpp:294-313 and the sphere upload :213-216.  The library re-packs them on the device
 * (see DESIGN.md "HBM layout"); the caller keeps ownership of everything it passes. */
GPT_API int gpt_upload_scene(gpt_ctx* ctx,
                             const void* nodes, size_t n_node_vec4,
                             const void* tris, size_t n_tri_vec4,
                             const int32_t* tri_indices, size_t n_indices,
                             const gpt_sphere* spheres, size_t n_spheres,
                             const gpt_box* boxes, size_t n_boxes,
                             int on_device);

/* ---- per-triangle materials ----
 * The reference shades every triangle with ONE global material (info.col / emi / triCurrentMat, tracer.cu:131-135;
 * tinyobj's .mtl data is dropped, utilfun.cpp:457-497).  With a table set, a triangle hit takes colour, emission
 * and Mat from materials[face_material[f]], f = the original OBJ face index the reference's index array maps the
 * hit to (gpuTriIndices[hitIndex], cudaUtils.h:452-456); faces outside the table keep the global material.
 * Host or device pointers; copied.  n_faces = 0 restores the reference behaviour. */
typedef struct { float col[3]; float emi[3]; int32_t mat; } gpt_tri_material;   /* 28 B */
GPT_API int gpt_set_triangle_materials(gpt_ctx* ctx, const gpt_tri_material* materials, size_t n_materials,
                                       const int32_t* face_material, size_t n_faces);

/* ---- device-side BVH builder ----
 * Replaces, for scenes where the single-threaded host build is the bottleneck (2.9 s per 100 k triangles), the
 * pipeline  BVH myBVH(&scene, Platform(), BuildParams()); new CudaBVH(myBVH, BVHLayout_Compact)  of
 * BasicScene.cpp:283-292 (SplitBVHBuilder.cpp:32-164, CudaBVH.cpp:121-270).  Input: the indexed mesh
 * uf::loadIndexedTris returns (utilfun.cpp:454-524; verts = n_verts*3 floats, tris = n_tris*3 ints), host
 * (on_device=0) or device pointers.  Output: DEVICE arrays in exactly the layout CudaBVH::getGpuNodes /
 * getDebugTri / getGpuTriIndices hand out, so `out->bvh` can be stored into kernelInfo.bvhData or passed to
 * gpt_upload_scene(..., on_device=1).  The tree is built by PLOC clustering over the Morton-sorted triangles (area-driven
 * merges, deterministic) with a SAH leaf collapse at the reference's cost model 1/1 -- not the reference's SBVH (no spatial
 * splits): hit distances are identical, tie-breaking between triangles with bit-equal t may differ.  GPT_FLAG_BUILD_LBVH asks
 * for the plain Morton radix tree instead.  The arrays belong to the caller until gpt_free_bvh. */
typedef struct {
    gpt_bvh_data bvh;
    uint32_t n_inner_nodes, n_leaves, n_triangles, depth;
    float build_ms;                 /* device + host time of the call */
} gpt_built_bvh;
GPT_API int gpt_build_bvh(gpt_ctx* ctx, const float* verts, size_t n_verts, const int32_t* tris, size_t n_tris,
                          int on_device, uint32_t max_leaf_size, gpt_built_bvh* out);
GPT_API int gpt_free_bvh(gpt_ctx* ctx, gpt_built_bvh* bvh);
/* plain cudaMemcpy device -> host (so C callers and the ctypes tests can inspect built arrays without the runtime) */
GPT_API int gpt_copy_to_host(gpt_ctx* ctx, void* dst_host, const void* src_device, size_t bytes);

/* ---- render launch ---- */
/* Drop-in for launchKernel (tracer.cu:405-415): ONE progressive pass (1 spp) with the reference's
 * exact field meanings: vec3 running mean with per-frame clamp into info->accumBuffer, RGBA8 pack into
 * info->dev_drawRes (tracer.cu:386-398), seed = info->hash.  If the BVH device pointers in `info` differ from
 * the ones packed last (address or element count of any of the three arrays) they are (re-)packed first; spheres are read
 * live from info->sphereTex on every launch, like the reference does (tracer.cu:39).
 * POINTER-IDENTITY RULE: the reference kernel reads dev_triNode / dev_triPtr / dev_triIndicesPtr live on every launch; this
 * library traverses a re-packed copy and cannot see writes INTO those arrays.  A caller that rewrites them in place (refit,
 * animation), or frees and re-allocates them and may get the same addresses back, must call gpt_invalidate_scene() before the
 * next gpt_launch; the INTEGRATION.md adapter does so wherever the reference re-uploads.  Asynchronous on `stream` (see the
 * header comment for the calls that may block). */
GPT_API int gpt_launch(gpt_ctx* ctx, const gpt_kernel_info* info, gpt_stream stream);
/* forget which kernelInfo BVH arrays the packed tree was made from: the next gpt_launch carrying BVH pointers re-packs them */
GPT_API int gpt_invalidate_scene(gpt_ctx* ctx);

/* Native contract: adds `sample_count` samples with GLOBAL indices [sample_begin, sample_begin+count)
 * to accum_f4 (device, W*H float4 = sum r,g,b + sample count; never clamped).  RNG is keyed by
 * (seed, pixel, global sample, bounce), so any split of the index range over GPUs gives the same sum.
 * info->accumBuffer / dev_drawRes / hash / constantPdf are ignored here. */
GPT_API int gpt_render(gpt_ctx* ctx, const gpt_kernel_info* info, float* accum_f4,
                       uint64_t seed, uint64_t sample_begin, uint32_t sample_count, gpt_stream stream);

/* gpt_render without any allocation.  NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same contract, same
 * samples and bit-identical sums as gpt_render on the wavefront pipeline, but the pipeline's scratch (per-path state of one wave,
 * per-pixel sums, queue counters) is a fixed file-scope device array inside the library instead of buffers grown on demand: 30 MiB
 * that EVERY GPU that loads the library holds until the process exits.  Limits, checked on entry (GPT_ERR_INVALID and a message
 * from gpt_last_error_string beyond them; use gpt_render there):
 *     info->width * info->height <= GPT_STATIC_MAX_PIXELS (262 144, e.g. 512 x 512);  info->depth <= GPT_STATIC_MAX_DEPTH (64);
 *     any sample_count: calls of more than GPT_STATIC_MAX_PATHS (262 144) paths run as several waves.
 * The scene must have been uploaded before (gpt_upload_scene / gpt_set_triangle_materials are the set-up calls and do allocate);
 * BVH pointers in info->bvhData are ignored (packing them allocates), info->sphereTex is read live as in gpt_render.  The tree
 * in use is traversed (gpt_tree_info: as uploaded until a trial of earlier gpt_render / gpt_launch calls has chosen another; the
 * 4-wide tree only if such a call already built it); the call itself takes part in no timed trial, no ray
 * re-ordering, no per-kernel timing.  The call itself makes no cudaMalloc / cudaFree and never synchronises; the first call on
 * each device resolves the array's device address (cudaGetSymbolAddress), so make one call before capturing.
 * SHARED STATE: all calls on a device share the array, whichever context they come through.  It keeps the last call's contents;
 * what must start at zero is cleared on `stream` at the start of every call.  Calls that overlap on the GPU (two streams, or two
 * host threads) race on it: serialise them on one stream. */
#define GPT_STATIC_MAX_PIXELS (256 * 1024)
#define GPT_STATIC_MAX_PATHS  (256 * 1024)
#define GPT_STATIC_MAX_DEPTH  64
GPT_API int gpt_render_static(gpt_ctx* ctx, const gpt_kernel_info* info, float* accum_f4,
                              uint64_t seed, uint64_t sample_begin, uint32_t sample_count, gpt_stream stream);

/* mean = sum/count, clamp [0,1], pack like tracer.cu:394-398.  Either output may be NULL. */
GPT_API int gpt_resolve(gpt_ctx* ctx, const float* accum_f4, int width, int height,
                        uint32_t* out_rgba8, float* out_rgb_f32, gpt_stream stream);

/* End-to-end convenience with HOST output: zero an internal float4 buffer, render `spp` samples,
 * resolve, copy RGBA8 (and optionally the float4 sums) back to host memory, synchronise. */
GPT_API int gpt_render_frame_host(gpt_ctx* ctx, const gpt_kernel_info* info, uint64_t seed,
                                  uint64_t sample_begin, uint32_t spp,
                                  uint32_t* host_rgba8, float* host_accum_f4);

/* ---- multi-GPU: one process per GPU, peer-memory frame reduce fused with accumulate + resolve ----
 * The reference is single-GPU (utilfun.cpp:73,95).  Here every rank renders a disjoint slice of a progressive frame's
 * global sample indices into its own per-frame float4 buffer; instead of ncclReduce + add + resolve, the ROOT reads the
 * peers' buffers directly over NVLink (CUDA IPC mappings) inside ONE kernel that sums the ranks in rank order, adds
 * the frame into the running sum and resolves the pixel.
 *   gpt_frame_alloc      library-owned frame buffers (n_buffers x W*H float4, one allocation) + the IPC handle of the
 *                        allocation; the caller exchanges handles between ranks (torch.distributed all_gather, MPI ...)
 *   gpt_frame_open_peers maps every other rank's allocation (handles[r], r != rank); world <= 16
 *   gpt_frame_ptr        this rank's buffer `index` (render target for gpt_render after gpt_frame_clear)
 *   gpt_reduce_resolve   root only: total += sum_r frame_r[index]; resolve into out_rgba8 / out_rgb_f32 (either NULL)
 * Ordering between ranks is the caller's: the peers' renders must have completed (a stream-ordered collective such
 * as a 4-byte all-reduce, or a host barrier after a sync) before the root launches gpt_reduce_resolve, and a buffer
 * must not be cleared again before the root has read it -- alternate between n_buffers >= 2. */
typedef struct { unsigned char reserved[64]; } gpt_ipc_mem_handle;      /* cudaIpcMemHandle_t */
GPT_API int gpt_frame_alloc(gpt_ctx* ctx, int width, int height, int n_buffers, gpt_ipc_mem_handle* handle_out);
GPT_API int gpt_frame_open_peers(gpt_ctx* ctx, const gpt_ipc_mem_handle* handles, int world, int rank);
GPT_API float* gpt_frame_ptr(gpt_ctx* ctx, int index);
GPT_API int gpt_frame_clear(gpt_ctx* ctx, int index, gpt_stream stream);
GPT_API int gpt_reduce_resolve(gpt_ctx* ctx, int index, float* total_accum_f4, uint32_t* out_rgba8, float* out_rgb_f32, gpt_stream stream);
/* the same kernel over explicit device pointers (frames[r], r = 0..n-1, all W*H float4): single-process use and tests */
GPT_API int gpt_reduce_resolve_ptrs(gpt_ctx* ctx, const float* const* frames, int n_frames, int width, int height,
                                    float* total_accum_f4, uint32_t* out_rgba8, float* out_rgb_f32, gpt_stream stream);
GPT_API int gpt_frame_free(gpt_ctx* ctx);

/* ---- test hooks / diagnostics ---- */
/* first bounce of every pixel: geometry type, remapped triangle index (or -1), sphere/box index
 * (or -1) and hit distance; device or NULL outputs of W*H elements. */
GPT_API int gpt_trace_primary(gpt_ctx* ctx, const gpt_kernel_info* info, uint64_t seed, uint64_t sample,
                              int32_t* geom, int32_t* tri_idx, int32_t* prim_idx, float* t, gpt_stream stream);
/* nearest triangle hit of arbitrary rays (device arrays: o,d = n*3 floats): remapped index,
 * slot address in the caller's triangle array (as the reference's hitIndex before remap), t */
GPT_API int gpt_trace_rays(gpt_ctx* ctx, const float* o, const float* d, size_t n, int cull_back_faces,
                           int32_t* tri_idx, int32_t* tri_addr, float* t, gpt_stream stream);
GPT_API int gpt_get_stats(gpt_ctx* ctx, gpt_stats* out);      /* synchronises the context's streams */
GPT_API int gpt_reset_stats(gpt_ctx* ctx);
/* The 4-wide tree in use (built lazily by the first launch that wants it): *n_wide = nodes (0 = binary tree in use),
 * *levels = wide levels.  Returns GPT_OK, or the error of the build. */
GPT_API int gpt_wide_info(gpt_ctx* ctx, int* n_wide, int* levels);
/* Ray reordering before the secondary extends (images are bit-identical with and without): *state = 1 on, 0 off,
 * -1 still on trial (scenes of >= 256 Ki packed triangles, when neither GPT_WF_SORT nor GPT_SORT_MIN_TRIS is set: part of the
 * trial gpt_trial_info describes); the two figures are the trial's whole-call times for the tree in use. */
GPT_API int gpt_reorder_info(gpt_ctx* ctx, int* state, double* ns_per_ray_path_order, double* ns_per_ray_reordered);
/* The rebuilt tree.  When arrays in the reference's layout are packed (gpt_upload_scene, or new BVH pointers in gpt_launch) and hold
 * 16 .. 4 Mi triangle references, the library also builds a tree of its own over the SAME references (device PLOC builder, the
 * caller's exact vertices; a few ms) and lets the per-scene trial decide between the two: the reference's SBVH is shaped for its own
 * kernel, and the extend kernel here runs 15-20 % faster on the rebuilt one.  Distances, face indices and slots are those of the
 * uploaded arrays bit for bit; what can differ is which of two triangles hit at bit-equal distance is met first (as for the 4-wide
 * tree).  Not built for arrays that come from gpt_build_bvh, nor under GPT_FLAG_KEEP_TREE / _BINARY_BVH / _WIDE_BVH / _MEGAKERNEL.
 * *state = 1 in use, 0 not, -1 on trial; *ns_per_ray = its timed trial calls (0 = not measured). */
GPT_API int gpt_rebuild_info(gpt_ctx* ctx, int* state, int* n_inner_nodes, int* depth, double* ns_per_ray, float* build_ms);
/* All candidate trees at once.  *tree_in_use: 0 the binary tree as uploaded, 1 its 4-wide collapse, 2 the rebuilt tree (above), 3 the
 * uploaded tree with its leaves split: every leaf of 3 .. 16 references replaced by a small SAH subtree (one thread per leaf at
 * pack time; the reference builder stops at leaves of up to 8 references, the kernels here want 1-2) -- what the reference builder
 * itself would have produced with a leaf cap of 2, and on large meshes a better tree than the rebuilt one.  Built and tried under the
 * same rules as the rebuilt tree (GPT_FLAG_KEEP_TREE; GPT_SPLIT=0 / 1 in the environment), with the same caveat (t-ties).
 * ns_per_ray[i] / n_inner_nodes[i]: the trial's figure (0 = not measured) and the node count (0 = not built) of tree i. */
GPT_API int gpt_tree_info(gpt_ctx* ctx, int* tree_in_use, int* pending, double ns_per_ray[4], int n_inner_nodes[4]);
/* The same for the single-launch megakernel (calls of fewer than GPT_SMALL_FRAME_PATHS paths, e.g. one launchKernel of a small
 * window): it keeps a trial of its own over the binary trees (0, 2, 3; calls of >= 16 Ki paths, three per candidate, the timed ones
 * synchronise) because what it prefers differs from the pipeline's choice.  ns_per_ray[1] stays 0. */
GPT_API int gpt_megakernel_tree_info(gpt_ctx* ctx, int* tree_in_use, int* pending, double ns_per_ray[4]);
/* The per-scene trial behind both choices (tree and ray order): *pending = 1 while large render calls (>= 1 Mi paths; >= 393 216 when a rebuilt tree is a candidate) are still being timed;
 * ns_per_ray = whole-call ns per ray of the timed calls so far: [0] binary tree, [1] 4-wide tree (both in path order),
 * [2] the faster tree with rays re-ordered; 0 = not measured (candidate forced, closed, or not reached yet). */
GPT_API int gpt_trial_info(gpt_ctx* ctx, int* pending, double ns_per_ray[3]);
GPT_API int gpt_get_kernel_times(gpt_ctx* ctx, gpt_kernel_times* out);  /* synchronises; see GPT_FLAG_TIME_KERNELS */
/* Peaks of THIS device measured by micro-benchmarks (about 0.2 s, needs 2 GiB of free device memory; synchronises): the
 * denominators bench.py's roofline fractions are stated against.  gather_*: GB/s of random 64-byte fetches (4 x LDG.128, the shape
 * of one BVH node fetch) from a working set resident in L1 (64 KiB) / in L2 (16 MiB) / only in HBM (2 GiB), issued from the extend
 * kernel's launch shape; warp_inst_per_s: warp instructions issued per second with all four schedulers of every SM busy;
 * stream_*: GB/s of coalesced 16-byte loads sweeping an L1- / L2-resident working set (the ceiling of that level). */
typedef struct { double gather_l1_gbs, gather_l2_gbs, gather_hbm_gbs, warp_inst_per_s, stream_l1_gbs, stream_l2_gbs; } gpt_peaks;
GPT_API int gpt_measure_peaks(gpt_ctx* ctx, gpt_peaks* out);

/* ---- host-side helpers mirroring the reference's scene setup (no GPU needed) ---- */
GPT_API uint64_t gpt_hash64(uint64_t key);                                   /* uf::hash, utilfun.cpp:380-389 */
GPT_API int  gpt_default_spheres(gpt_sphere* out, size_t cap, size_t* n);   /* BasicScene.cpp:177-218 */
GPT_API void gpt_default_kernel_info(gpt_kernel_info* info, int width, int height); /* BasicScene.cpp:220-259 */
GPT_API void gpt_set_camera(gpt_cam_info* cam, const float pos[3], float yaw_deg, float pitch_deg); /* BasicScene.cpp:42-64 */

/* ---- headless scene / image I/O (no GPU needed) ---- */
/* uf::loadIndexedTris (utilfun.cpp:454-524): indexed triangle mesh of an OBJ file (fan-triangulated, 0-based,
 * materials dropped); several shapes are merged (the reference asserts exactly one, :474).  The arrays are
 * malloc'd by the library; release with gpt_free_host. */
GPT_API int  gpt_load_obj(const char* path, float** verts, size_t* n_verts, int32_t** tris, size_t* n_tris);
/* OBJ plus its mtllib: additionally one material index per triangle (usemtl; -1 = none) and one
 * gpt_tri_material per material name (col = Kd, emi = Ke, Mat from illum: 0-2 DIFF, 3 METAL, 5 SPEC, 4/6/7/9 REFR),
 * ready for gpt_set_triangle_materials.  All four arrays are released with gpt_free_host. */
GPT_API int  gpt_load_obj_materials(const char* path, float** verts, size_t* n_verts, int32_t** tris, size_t* n_tris,
                                    int32_t** face_material, gpt_tri_material** materials, size_t* n_materials);
GPT_API void gpt_free_host(void* p);
/* RGBA8 frame as dev_drawRes holds it (0x00BBGGRR, row 0 = bottom; copied to the host by the caller) -> PNG, top
 * row first: replaces the GL texture blit + quad (BasicScene.cpp:424-432,468; renderer.cpp:224-245). */
GPT_API int  gpt_write_png(const char* path, const uint32_t* host_rgba8, int width, int height);
/* accumulation checkpoint / resume: host float4 sums + samples done + seed; because sample indices
 * are global, rendering [0,a) , saving, loading and rendering [a,b) equals rendering [0,b) in one go */
GPT_API int  gpt_save_accum(const char* path, const float* host_accum_f4, int width, int height, uint64_t samples_done, uint64_t seed);
GPT_API int  gpt_load_accum(const char* path, float* host_accum_f4, int width, int height, uint64_t* samples_done, uint64_t* seed);
/* linear float RGB (gpt_resolve's out_rgb_f32, copied to the host) -> PFM, bottom row first */
GPT_API int  gpt_write_pfm(const char* path, const float* host_rgb, int width, int height);
/* the same image as OpenEXR 2.0 (north_star item 3: headless PNG/EXR output instead of the GL texture, BasicScene.cpp:424-432):
 * single-part scan-line file, uncompressed, 32-bit FLOAT channels B, G, R, top row first */
GPT_API int  gpt_write_exr(const char* path, const float* host_rgb, int width, int height);

#ifdef __cplusplus
}
#endif

#ifdef __cplusplus
#include <cstddef>
static_assert(sizeof(gpt_sphere) == 44, "Sphere ABI");
static_assert(sizeof(gpt_tri_material) == 28, "gpt_tri_material");
static_assert(sizeof(gpt_box) == 52, "Box ABI");
static_assert(sizeof(gpt_cam_info) == 84, "CamInfo ABI");
static_assert(sizeof(gpt_bvh_data) == 48, "BVHData ABI");
static_assert(sizeof(gpt_kernel_info) == 296, "kernelInfo ABI");
static_assert(offsetof(gpt_kernel_info, accumBuffer) == 32, "kernelInfo.accumBuffer");
static_assert(offsetof(gpt_kernel_info, cam) == 48, "kernelInfo.cam");
static_assert(offsetof(gpt_kernel_info, sphereTex) == 152, "kernelInfo.sphereTex");
static_assert(offsetof(gpt_kernel_info, depth) == 172, "kernelInfo.depth");
static_assert(offsetof(gpt_kernel_info, constantPdf) == 176, "kernelInfo.constantPdf");
static_assert(offsetof(gpt_kernel_info, bvhData) == 184, "kernelInfo.bvhData");
static_assert(offsetof(gpt_kernel_info, triCurrentMat) == 244, "kernelInfo.triCurrentMat");
static_assert(offsetof(gpt_kernel_info, bkColor) == 252, "kernelInfo.bkColor");
static_assert(offsetof(gpt_kernel_info, emi) == 268, "kernelInfo.emi");
#endif
#endif /* GPT_H */
