// gpt_abi.cu -- implementation of the C-ABI declared in include/gpt.h.
//
// Host-side orchestration only: context, scene upload + device re-pack, parameter flattening and
// kernel launches.  No CPU fallback exists: without a usable CUDA device every compute entry fails
// with GPT_ERR_CUDA and a message.
#include "../../include/gpt.h"
#include "gpt_kernels.h"
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <mutex>
#include <new>
#include <string>

struct gpt_ctx {
    int device = 0;
    int sm_count = 0;
    uint32_t flags = 0;
    // uploaded scene (owned)
    float4* nodes = nullptr; float4* tris = nullptr; int n_nodes = 0; int n_tris = 0; int tree_depth = 0;
    float* spheres = nullptr; int n_spheres = 0;
    float* boxes = nullptr; int n_boxes = 0;
    float* tri_mats = nullptr; int* face_mat = nullptr; int n_tri_mats = 0, n_faces = 0;   // N4, survive scene re-uploads
    bool has_scene = false;
    // N3: 4-wide quantised tree over the packed binary one, built on first use (gpt_wide.cu)
    float4* wnodes = nullptr; int n_wide = 0, wide_levels = 0; bool wide_tried = false;
    // a second packed tree over the same triangle references, made by the device builder when a reference-layout tree is packed
    // (gptk_rebuild); own_tree = the uploaded arrays came from gpt_build_bvh of this context (nothing to gain from rebuilding)
    float4* alt_nodes = nullptr; float4* alt_tris = nullptr; int alt_n_nodes = 0, alt_n_tris = 0, alt_depth = 0; float alt_build_ms = 0.0f;
    // ... and the uploaded tree with its larger leaves split into small SAH subtrees (gptk_split_leaves)
    float4* spl_nodes = nullptr; float4* spl_tris = nullptr; int spl_n_nodes = 0, spl_depth = 0;
    bool own_tree = false; const void* built_nodes[8] = {}; int built_next = 0;
    // per-scene choice of traversal structure and ray order (traversal_plan): flags and environment win, else a timed trial
    int tune_step = 0; bool tune_done = false; int tuned_tree = 0; bool tuned_sort = false;      // tree ids: 0 uploaded binary, 1 its 4-wide collapse, 2 rebuilt, 3 leaves split
    // the single-launch megakernel (small calls) keeps its own trial over the binary trees: what it prefers differs from the pipeline
    int mega_step = 0; bool mega_done = false; int mega_tree = 0; double mega_ns[4] = {0.0, 0.0, 0.0, 0.0};
    double tune_ns[5] = {0.0, 0.0, 0.0, 0.0, 0.0};     // ns per ray: binary tree / wide tree in path order, chosen tree re-ordered, rebuilt tree, split-leaf tree (path order)
    // identity of the reference arrays the packed BVH was built from (gpt_launch re-packs on change)
    const void* src_nodes = nullptr; const void* src_tris = nullptr; const void* src_idx = nullptr;
    size_t src_n_nodes = 0, src_n_tris = 0, src_n_idx = 0;
    // scratch
    unsigned int* work_counter = nullptr;
    unsigned long long* stats = nullptr;       // [0]=rays [1]=paths [2..17]=rays by bounce
    float4* frame_accum = nullptr; uint32_t* frame_rgba = nullptr; size_t frame_pixels = 0;
    WaveBuffers wave = {}; size_t wave_pixels = 0; size_t wave_counters = 0;   // wavefront scratch
    WaveLaunchCfg wave_cfg = {};
    bool wave_dirty = false;                  // a render call failed half-way: the queue counters must be re-zeroed before the next
    unsigned long long launches = 0;
    // multi-GPU peer frames
    float4* frames = nullptr; int frame_w = 0, frame_h = 0, frame_n = 0;
    void* peer_base[16] = {}; int peer_world = 0, peer_rank = 0;
    KernelTimer timer; double kt_ms[KT_CLASSES] = {}; unsigned long long kt_n[KT_CLASSES] = {};
    std::string err;
};

static thread_local std::string g_create_err;

static int fail(gpt_ctx* c, int code, const char* what, cudaError_t e = cudaSuccess)
{
    char buf[512];
    if (e != cudaSuccess) snprintf(buf, sizeof(buf), "%s: %s (%s)", what, cudaGetErrorString(e), cudaGetErrorName(e));
    else snprintf(buf, sizeof(buf), "%s", what);
    if (c) c->err = buf; else g_create_err = buf;
    return code;
}
#define CU(c, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail((c), GPT_ERR_CUDA, #call, e_); } while (0)

extern "C" {

const char* gpt_version(void) { return "gpt_b200 0.1 (sm_100a)"; }

const char* gpt_last_error_string(gpt_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

int gpt_create(gpt_ctx** out, int device)
{
    if (!out) return fail(nullptr, GPT_ERR_INVALID, "gpt_create: out is NULL");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) return fail(nullptr, GPT_ERR_CUDA, "gpt_create: no CUDA device (this library has no CPU path)", e);
    if (device < 0 || device >= n) return fail(nullptr, GPT_ERR_INVALID, "gpt_create: device index out of range");
    CU(nullptr, cudaSetDevice(device));
    gpt_ctx* c = new (std::nothrow) gpt_ctx();
    if (!c) return fail(nullptr, GPT_ERR_INVALID, "gpt_create: out of host memory");
    c->device = device;
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess) { delete c; return fail(nullptr, GPT_ERR_CUDA, "cudaGetDeviceProperties", e); }
    c->sm_count = prop.multiProcessorCount;
    if ((e = cudaMalloc(&c->work_counter, sizeof(unsigned int))) != cudaSuccess ||
        (e = cudaMalloc(&c->stats, 32 * sizeof(unsigned long long))) != cudaSuccess ||
        (e = cudaMemset(c->stats, 0, 32 * sizeof(unsigned long long))) != cudaSuccess) {
        cudaFree(c->work_counter); cudaFree(c->stats); delete c;
        return fail(nullptr, GPT_ERR_CUDA, "gpt_create: scratch allocation", e);
    }
    *out = c;
    return GPT_OK;
}

static void free_wide(gpt_ctx* c)
{
    cudaFree(c->wnodes); c->wnodes = nullptr; c->n_wide = 0; c->wide_levels = 0; c->wide_tried = false;
    c->tune_step = 0; c->tune_done = false; c->tuned_tree = 0; c->tuned_sort = false;      // a new tree: the questions are open again
    c->tune_ns[0] = c->tune_ns[1] = c->tune_ns[2] = c->tune_ns[3] = c->tune_ns[4] = 0.0;
    c->mega_step = 0; c->mega_done = false; c->mega_tree = 0; c->mega_ns[0] = c->mega_ns[1] = c->mega_ns[2] = c->mega_ns[3] = 0.0;
}

static void free_alt(gpt_ctx* c)
{
    cudaFree(c->alt_nodes); cudaFree(c->alt_tris); c->alt_nodes = c->alt_tris = nullptr;
    c->alt_n_nodes = c->alt_n_tris = c->alt_depth = 0; c->alt_build_ms = 0.0f;
    cudaFree(c->spl_nodes); cudaFree(c->spl_tris); c->spl_nodes = c->spl_tris = nullptr; c->spl_n_nodes = c->spl_depth = 0;
}

static void free_scene(gpt_ctx* c)
{
    free_wide(c); free_alt(c);
    cudaFree(c->nodes); cudaFree(c->tris); cudaFree(c->spheres); cudaFree(c->boxes);
    c->nodes = c->tris = nullptr; c->spheres = c->boxes = nullptr;
    c->n_nodes = c->n_tris = c->n_spheres = c->n_boxes = 0;
    c->has_scene = false;
    c->src_nodes = c->src_tris = c->src_idx = nullptr;
}

static void free_wave(gpt_ctx* c)
{
    WaveBuffers& w = c->wave;
    cudaFree(w.ray); cudaFree(w.mask_z); cudaFree(w.accu); cudaFree(w.lsum); cudaFree(w.hit);
    cudaFree(w.queue[0]); cudaFree(w.queue[1]); cudaFree(w.counters);
    cudaFree(w.sort_keys[0]); cudaFree(w.sort_keys[1]); cudaFree(w.sort_vals); cudaFree(w.sort_tmp);
    w = WaveBuffers{}; c->wave_pixels = 0; c->wave_counters = 0;
}

// wavefront scratch: per-path state for one wave (104 B/path) + one launch sum per pixel.  A wave holds as many
// samples of the frame as fit GPT_WAVE_MAX_PATHS (5 GB of the 180 GB HBM at the cap); grown on demand.
#define GPT_WAVE_MAX_PATHS ((size_t)64 << 20)
static int ensure_wave(gpt_ctx* c, size_t n_pix, unsigned int depth, uint32_t sample_count)
{
    const size_t nc = gptk_wave_counter_count(depth);
    size_t want = n_pix * (size_t)(sample_count ? sample_count : 1);
    if (want > GPT_WAVE_MAX_PATHS) want = n_pix > GPT_WAVE_MAX_PATHS ? GPT_WAVE_MAX_PATHS : (GPT_WAVE_MAX_PATHS / n_pix) * n_pix;
    if (c->wave_pixels >= n_pix && c->wave_counters >= nc && c->wave.capacity >= want) return GPT_OK;
    CU(c, cudaDeviceSynchronize());
    if (want < c->wave.capacity) want = c->wave.capacity;
    free_wave(c);
    WaveBuffers& w = c->wave;
    CU(c, cudaMalloc(&w.ray, 2 * 2 * want * sizeof(float4)));        // double-buffered by bounce parity, like the queues
    CU(c, cudaMalloc(&w.mask_z, 2 * want * sizeof(float)));  CU(c, cudaMalloc(&w.accu, want * sizeof(float4)));
    CU(c, cudaMalloc(&w.hit, want * sizeof(int2)));
    CU(c, cudaMalloc(&w.queue[0], want * sizeof(uint32_t))); CU(c, cudaMalloc(&w.queue[1], want * sizeof(uint32_t)));
    CU(c, cudaMalloc(&w.lsum, n_pix * sizeof(float4)));
    CU(c, cudaMalloc(&w.counters, nc * sizeof(unsigned int)));
    CU(c, cudaMemset(w.counters, 0, nc * sizeof(unsigned int)));       // kept zero between calls by the pipeline's last kernel
    w.capacity = want;
    c->wave_pixels = n_pix; c->wave_counters = nc;
    w.sort_keys[0] = w.sort_keys[1] = w.sort_vals = nullptr; w.sort_tmp = nullptr; w.sort_tmp_bytes = 0;
    return GPT_OK;
}

// ---- which tree, which ray order -------------------------------------------------------------------------------------------
// Two choices change nothing in the image and a lot in the frame time, and neither follows from the triangle count alone
// (profiles/r01_schedules.md #15-#25): the 4-wide quantised tree (gpt_wide.cu) against the binary tree as uploaded -- -10 % on a
// 100 k-triangle dragon in a box, +3 % on 250 k triangles that fill space, +40 % at 20 M -- and re-ordering the secondary rays
// before each extend (gpt_wavefront.cu, ~40 ps per ray) -- +11 % for the binary tree on 2 M space-filling triangles, -15 % on
// the 1.6 M-triangle dragon.  Policy, per dimension:
//   tree : GPT_FLAG_BINARY_BVH / GPT_FLAG_WIDE_BVH force it; GPT_WIDE_MIN_TRIS=n sets a size threshold; the megakernel and
//          scenes under 64 Ki packed triangles use the binary tree; otherwise it is put on trial (default until decided:
//          wide from 1 Mi triangles);
//   order: GPT_WF_SORT=0 never, =1..9 always (key layout); GPT_SORT_MIN_TRIS=n a size threshold; never under 256 Ki packed
//          triangles; otherwise on trial (default until decided: path order).
// Trial: the first large render calls (>= 1 Mi paths) after a scene upload run each open candidate three times -- one warm-up
// call, two timed calls bracketed by events and a stream synchronise, the faster of which counts (one noisy measurement must
// not pin the slower path for the lifetime of the scene) -- binary / path order, wide / path order, then the faster tree
// re-ordered; ns per ray decide (the wide tree must win by 2 %, re-ordering by 3 %).  At most six calls synchronise.
// GPT_FLAG_NO_TRIAL (or GPT_NO_TRIAL=1): never -- every render call is then asynchronous and capturable; the size defaults apply.
static const bool k_env_no_trial = []{ const char* e = getenv("GPT_NO_TRIAL"); return e && *e && *e != '0'; }();
static const long long k_env_wide_min = []{ const char* e = getenv("GPT_WIDE_MIN_TRIS"); return e ? atoll(e) : -2ll; }();   // -2 = unset
static const int k_env_sort = []{ const char* e = getenv("GPT_WF_SORT"); return e ? atoi(e) : -1; }();
static const long long k_env_sort_min = []{ const char* e = getenv("GPT_SORT_MIN_TRIS"); return e ? atoll(e) : -2ll; }();
// The rebuilt tree (gptk_rebuild): GPT_REBUILD=0 never builds it, =1 always traverses it (where one could be built), unset: built for
// reference-layout trees of 1 Ki .. 4 Mi packed triangles that do not come from gpt_build_bvh, and put on trial with the others.
// GPT_FLAG_KEEP_TREE, GPT_FLAG_BINARY_BVH, GPT_FLAG_WIDE_BVH and GPT_FLAG_MEGAKERNEL all mean "the tree as uploaded".
static const int k_env_rebuild = []{ const char* e = getenv("GPT_REBUILD"); return (e && *e) ? atoi(e) : -1; }();
// The split-leaf tree (gptk_split_leaves) follows the same rules; GPT_SPLIT=0 / 1 switches it alone (1: always traversed, and
// GPT_REBUILD=1 wins over it); GPT_SPLIT_MAX_LEAF sets the leaf cap (2)
static const int k_env_split = []{ const char* e = getenv("GPT_SPLIT"); return (e && *e) ? atoi(e) : -1; }();

// Which pipeline: the wavefront pipeline needs ~19 launches per call and each costs ~19 us of launch gap + drain, so below a few
// hundred thousand paths per call the single-launch megakernel is faster (cornell 512x512, 1 spp, depth 4: 0.17 ms against
// 0.20 ms, and against 0.21 ms for the reference's kernel; profiles/r01_small_frames.jsonl).  Same samples, same images.
static const long long k_env_small_frame = []{ const char* e = getenv("GPT_SMALL_FRAME_PATHS"); return (e && *e) ? atoll(e) : 393216ll; }();
static bool use_megakernel(const gpt_ctx* c, size_t n_paths)
{
    if (c->flags & GPT_FLAG_MEGAKERNEL) return true;
    if (c->flags & (GPT_FLAG_WAVEFRONT | GPT_FLAG_FAST)) return false;      // the fast arithmetic exists in the pipeline's shade kernel only
    return (long long)n_paths < k_env_small_frame;
}

// 0 = binary, 1 = wide, -1 = open (trial)
static int tree_choice(const gpt_ctx* c)
{
    if (!c->nodes || c->n_nodes <= 0 || (c->flags & (GPT_FLAG_BINARY_BVH | GPT_FLAG_MEGAKERNEL))) return 0;
    if (c->wide_tried && !c->wnodes) return 0;                    // the tree could not be quantised: binary, same results
    if (c->flags & GPT_FLAG_WIDE_BVH) return 1;
    if (k_env_wide_min != -2) return k_env_wide_min >= 0 && (long long)c->n_tris >= k_env_wide_min ? 1 : 0;
    if (c->n_tris < (64 << 10)) return 0;
    return -1;
}
static int sort_choice(const gpt_ctx* c)
{
    if (!c->nodes || (c->flags & GPT_FLAG_MEGAKERNEL) || k_env_sort == 0) return 0;
    if (k_env_sort > 0) return 1;
    if (k_env_sort_min != -2) return k_env_sort_min >= 0 && (long long)c->n_tris >= k_env_sort_min ? 1 : 0;
    if (c->n_tris < (256 << 10)) return 0;
    return -1;
}
// the rebuilt tree: 0 = not available / not allowed, 1 = forced, -1 = open (trial)
static int alt_choice(const gpt_ctx* c)
{
    if (!c->alt_nodes || (c->flags & (GPT_FLAG_KEEP_TREE | GPT_FLAG_BINARY_BVH | GPT_FLAG_WIDE_BVH | GPT_FLAG_MEGAKERNEL)) || k_env_rebuild == 0) return 0;
    return k_env_rebuild > 0 ? 1 : -1;
}
static int split_choice(const gpt_ctx* c)
{
    if (!c->spl_nodes || (c->flags & (GPT_FLAG_KEEP_TREE | GPT_FLAG_BINARY_BVH | GPT_FLAG_WIDE_BVH | GPT_FLAG_MEGAKERNEL)) || k_env_rebuild == 0 || k_env_split == 0) return 0;
    return k_env_split > 0 ? 1 : -1;
}
// the trees a scene may be traversed on (ids: 0 uploaded binary, 1 its 4-wide collapse, 2 rebuilt, 3 leaves split), in trial order
static int tree_cands(const gpt_ctx* c, int* ids)
{
    const int t = tree_choice(c), a = alt_choice(c), sp = split_choice(c);
    int n = 0;
    if (a == 1) { ids[n++] = 2; return n; }
    if (sp == 1) { ids[n++] = 3; return n; }
    if (t != 1) ids[n++] = 0;
    if (t != 0) ids[n++] = 1;
    if (a < 0) ids[n++] = 2;
    if (sp < 0) ids[n++] = 3;
    return n;
}
static double tree_ns(const gpt_ctx* c, int id) { return c->tune_ns[id >= 2 ? id + 1 : id]; }
// the fastest measured candidate; a later candidate must beat the best so far by 2 %
static int best_tree(const gpt_ctx* c, const int* ids, int n)
{
    int best = ids[0];
    for (int k = 1; k < n; ++k) {
        const double m = tree_ns(c, ids[k]), b = tree_ns(c, best);
        if (m > 0.0 && b > 0.0 && m < 0.98 * b) best = ids[k];
    }
    return best;
}
// what to use outside a trial call: forced, decided, or the default (the wide tree from 1 Mi triangles, else the tree as uploaded)
static int current_tree(const gpt_ctx* c)
{
    int ids[4];
    const int n = tree_cands(c, ids);
    if (n == 1) return ids[0];
    if (c->tune_done) return c->tuned_tree;
    if (tree_choice(c) < 0 && c->n_tris >= (1 << 20)) return 1;
    return ids[0];
}
static bool current_wide(const gpt_ctx* c) { return current_tree(c) == 1; }
static bool current_sort(const gpt_ctx* c)
{
    const int s = sort_choice(c);
    return s >= 0 ? s == 1 : (c->tune_done && c->tuned_sort);
}

struct TraversalPlan { int tree; bool sort; bool trial; bool timed; int slot; };     // slot = tune_ns index of a timed call
static TraversalPlan traversal_plan(const gpt_ctx* c, size_t n_paths)
{
    TraversalPlan pl = {current_tree(c), current_sort(c), false, false, -1};
    int ids[4];
    const int n_tree = tree_cands(c, ids), s = sort_choice(c);
    const int n_cand = n_tree + (s < 0 ? 1 : 0);
    // calls large enough to time: 1 Mi paths; with a rebuilt tree among the candidates every call the wavefront pipeline takes by
    // default (one 1-spp launchKernel of a 720p window is 0.9 M paths and must not be left on the slower tree for ever)
    const size_t min_paths = alt_choice(c) < 0 ? (size_t)393216 : ((size_t)1 << 20);
    if (c->tune_done || n_cand < 2 || n_paths < min_paths || (c->flags & GPT_FLAG_NO_TRIAL) || k_env_no_trial) return pl;
    const int cand = c->tune_step / 3;
    pl.trial = true; pl.timed = (c->tune_step % 3) != 0;
    if (cand < n_tree) {                                          // a tree candidate, rays in path order
        pl.tree = ids[cand];
        pl.sort = false;
        pl.slot = pl.tree >= 2 ? pl.tree + 1 : pl.tree;
    } else {                                                      // the fastest tree, re-ordered
        pl.tree = best_tree(c, ids, n_tree);
        pl.sort = true;
        pl.slot = 2;
    }
    return pl;
}
static void trial_advance(gpt_ctx* c)
{
    int ids[4];
    const int n_tree = tree_cands(c, ids), s = sort_choice(c);
    const int n_cand = n_tree + (s < 0 ? 1 : 0);
    if (++c->tune_step < 3 * n_cand) return;
    c->tune_done = true;
    c->tuned_tree = best_tree(c, ids, n_tree);
    const double path = tree_ns(c, c->tuned_tree);
    c->tuned_sort = s < 0 && c->tune_ns[2] > 0.0 && path > 0.0 && c->tune_ns[2] < 0.97 * path;
}
// point a DevScene at one of the context's trees (1 without a built wide tree falls back to the binary tree it was collapsed from)
static void set_tree(const gpt_ctx* c, DevScene* sc, int tree)
{
    const bool alt = tree == 2 && c->alt_nodes, spl = tree == 3 && c->spl_nodes;
    sc->nodes = alt ? c->alt_nodes : (spl ? c->spl_nodes : c->nodes); sc->tris = alt ? c->alt_tris : (spl ? c->spl_tris : c->tris);
    sc->n_nodes = alt ? c->alt_n_nodes : (spl ? c->spl_n_nodes : c->n_nodes); sc->n_tris = alt ? c->alt_n_tris : c->n_tris;
    sc->wnodes = (tree == 1) ? c->wnodes : nullptr;
}

static int ensure_sort(gpt_ctx* c, bool want)
{
    WaveBuffers& w = c->wave;
    if (!want || w.sort_vals != nullptr || w.capacity == 0) return GPT_OK;       // buffers are kept once allocated
    CU(c, cudaDeviceSynchronize());
    w.sort_tmp_bytes = gptk_wave_sort_tmp_bytes(w.capacity);
    CU(c, cudaMalloc(&w.sort_keys[0], w.capacity * sizeof(uint32_t))); CU(c, cudaMalloc(&w.sort_keys[1], w.capacity * sizeof(uint32_t)));
    CU(c, cudaMalloc(&w.sort_vals, w.capacity * sizeof(uint32_t)));
    CU(c, cudaMalloc(&w.sort_tmp, w.sort_tmp_bytes ? w.sort_tmp_bytes : 16));
    return GPT_OK;
}

// ---- the same question for the megakernel (calls of fewer than GPT_SMALL_FRAME_PATHS paths: one launchKernel of a small window) ----
// Candidates: the binary trees among tree_cands (no 4-wide traversal in the megakernel, no ray order to choose).  Calls of >= 16 Ki
// paths take part, three per candidate as above; whole-call ns per ray decide, a later candidate must win by 2 %.  Measured
// (profiles/r02_schedules.md #70): the megakernel prefers the split-leaf tree on every scene tried, also where the pipeline prefers
// the rebuilt one, hence a trial of its own.
static int mega_cands(const gpt_ctx* c, int* ids)
{
    int all[4];
    const int n_all = tree_cands(c, all);
    int n = 0;
    for (int k = 0; k < n_all; ++k) if (all[k] != 1) ids[n++] = all[k];
    if (n == 0) ids[n++] = 0;
    return n;
}
static int mega_best(const gpt_ctx* c, const int* ids, int n)
{
    int best = ids[0];
    for (int k = 1; k < n; ++k) {
        const double m = c->mega_ns[ids[k]], b = c->mega_ns[best];
        if (m > 0.0 && b > 0.0 && m < 0.98 * b) best = ids[k];
    }
    return best;
}
// one megakernel render call (both contracts) on the tree that policy picks
static int render_mega(gpt_ctx* c, const DevScene& sc_in, const RenderParams& p, float4* accum_f4, float* accum_v3, uint32_t* draw,
                       unsigned long long constant_pdf, unsigned long long sample_begin, uint32_t sample_count, cudaStream_t st)
{
    int ids[4];
    const int n = mega_cands(c, ids);
    const size_t n_paths = (size_t)p.width * p.height * sample_count;
    int tree = n == 1 ? ids[0] : (c->mega_done ? c->mega_tree : ((c->tune_done && c->tuned_tree != 1) ? c->tuned_tree : ids[0]));
    const bool trial = n > 1 && !c->mega_done && n_paths >= 16384 && !(c->flags & GPT_FLAG_NO_TRIAL) && !k_env_no_trial;
    const bool timed = trial && (c->mega_step % 3) != 0;
    if (trial) tree = ids[c->mega_step / 3];
    DevScene sc = sc_in;
    set_tree(c, &sc, tree);
    sc.wnodes = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    unsigned long long rays0 = 0, rays1 = 0;
    if (timed) {
        CU(c, cudaStreamSynchronize(st));
        CU(c, cudaMemcpy(&rays0, c->stats, sizeof(rays0), cudaMemcpyDeviceToHost));
        CU(c, cudaEventCreate(&e0)); CU(c, cudaEventCreate(&e1));
        CU(c, cudaEventRecord(e0, st));
    }
    cudaError_t e = accum_v3 ? gptk_render_legacy(sc, p, accum_v3, draw, constant_pdf, c->work_counter, c->stats, c->sm_count, st)
                             : gptk_render_mega(sc, p, accum_f4, sample_begin, sample_count, c->work_counter, c->stats, c->sm_count, st);
    c->launches += 1;
    if (e != cudaSuccess) { if (e0) { cudaEventDestroy(e0); cudaEventDestroy(e1); } return fail(c, GPT_ERR_CUDA, "render launch", e); }
    if (timed) {
        float ms = 0.0f;
        cudaEventRecord(e1, st);
        e = cudaEventSynchronize(e1);
        if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, e0, e1);
        if (e == cudaSuccess) e = cudaMemcpy(&rays1, c->stats, sizeof(rays1), cudaMemcpyDeviceToHost);
        cudaEventDestroy(e0); cudaEventDestroy(e1);
        if (e != cudaSuccess) return fail(c, GPT_ERR_CUDA, "megakernel tree trial", e);
        const double ns = rays1 > rays0 ? (double)ms * 1e6 / (double)(rays1 - rays0) : 0.0;
        if (ns > 0.0 && (c->mega_ns[tree] <= 0.0 || ns < c->mega_ns[tree])) c->mega_ns[tree] = ns;
    }
    if (trial && ++c->mega_step >= 3 * n) { c->mega_done = true; c->mega_tree = mega_best(c, ids, n); }
    return GPT_OK;
}

// one wavefront render call under that policy (both contracts)
static int render_wavefront(gpt_ctx* c, const DevScene& sc_in, const RenderParams& p, float4* accum_f4, float* accum_v3, uint32_t* draw,
                            unsigned long long constant_pdf, unsigned long long sample_begin, uint32_t sample_count, cudaStream_t st)
{
    int rc = ensure_wave(c, (size_t)p.width * p.height, p.depth, sample_count); if (rc) return rc;
    const TraversalPlan pl = traversal_plan(c, (size_t)p.width * p.height * sample_count);
    rc = ensure_sort(c, pl.sort); if (rc) return rc;
    DevScene sc = sc_in;
    set_tree(c, &sc, pl.tree);
    WaveBuffers wb = c->wave;
    if (!pl.sort) wb.sort_vals = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    unsigned long long rays0 = 0, rays1 = 0;
    if (pl.timed) {
        CU(c, cudaStreamSynchronize(st));
        CU(c, cudaMemcpy(&rays0, c->stats, sizeof(rays0), cudaMemcpyDeviceToHost));
        CU(c, cudaEventCreate(&e0)); CU(c, cudaEventCreate(&e1));
        CU(c, cudaEventRecord(e0, st));
    }
    if (!c->wave_cfg.ready) gptk_wave_launch_cfg(&c->wave_cfg, c->sm_count);
    if (c->wave_dirty) { CU(c, cudaMemsetAsync(wb.counters, 0, c->wave_counters * sizeof(unsigned int), st)); c->wave_dirty = false; }
    cudaError_t e = gptk_render_wavefront(sc, p, wb, c->wave_cfg, accum_f4, accum_v3, draw, constant_pdf, sample_begin, sample_count,
                                          c->stats, st, &c->launches, (c->flags & GPT_FLAG_TIME_KERNELS) ? &c->timer : nullptr);
    if (e != cudaSuccess) { c->wave_dirty = true; if (e0) { cudaEventDestroy(e0); cudaEventDestroy(e1); } return fail(c, GPT_ERR_CUDA, "render launch", e); }
    if (pl.timed) {
        float ms = 0.0f;
        cudaEventRecord(e1, st);
        e = cudaEventSynchronize(e1);
        if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, e0, e1);
        if (e == cudaSuccess) e = cudaMemcpy(&rays1, c->stats, sizeof(rays1), cudaMemcpyDeviceToHost);
        cudaEventDestroy(e0); cudaEventDestroy(e1);
        if (e != cudaSuccess) return fail(c, GPT_ERR_CUDA, "traversal trial", e);
        const double ns = rays1 > rays0 ? (double)ms * 1e6 / (double)(rays1 - rays0) : 0.0;
        if (ns > 0.0 && (c->tune_ns[pl.slot] <= 0.0 || ns < c->tune_ns[pl.slot])) c->tune_ns[pl.slot] = ns;      // best of the two timed calls
    }
    if (pl.trial) trial_advance(c);
    return GPT_OK;
}

void gpt_destroy(gpt_ctx* c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    free_scene(c);
    cudaFree(c->work_counter); cudaFree(c->stats); cudaFree(c->frame_accum); cudaFree(c->frame_rgba);
    cudaFree(c->tri_mats); cudaFree(c->face_mat);
    free_wave(c);
    gpt_frame_free(c);
    c->timer.release();
    delete c;
}

int gpt_set_flags(gpt_ctx* c, uint32_t flags)
{
    if (!c) return GPT_ERR_INVALID;
    c->flags = flags;
    return GPT_OK;
}

int gpt_sync(gpt_ctx* c, gpt_stream stream)
{
    if (!c) return GPT_ERR_INVALID;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize((cudaStream_t)stream));
    return GPT_OK;
}

// (re)build the packed BVH from reference-layout arrays that are already on the device
static int pack_bvh_from_device(gpt_ctx* c, const void* nodes, size_t n_node_vec4, const void* tris, size_t n_tri_vec4,
                                const int32_t* idx, size_t n_idx)
{
    cudaFree(c->nodes); cudaFree(c->tris); c->nodes = c->tris = nullptr; c->n_nodes = c->n_tris = 0;
    free_wide(c); free_alt(c);
    c->own_tree = false;
    for (const void* b : c->built_nodes) if (b && b == nodes) c->own_tree = true;
    if (n_node_vec4 == 0 || !nodes) return GPT_OK;     // sphere-only scene
    if (!tris || !idx) return fail(c, GPT_ERR_INVALID, "BVH nodes given without triangle / index arrays");
    RepackResult r;
    cudaError_t e = gptk_repack((const float4*)nodes, n_node_vec4, (const float4*)tris, n_tri_vec4, idx, n_idx, &r, 0);
    c->launches += 2;
    if (e != cudaSuccess) return fail(c, GPT_ERR_CUDA, "device re-pack", e);
    if (r.bad_layout) {
        char buf[160];
        if (r.bad_layout == 5) snprintf(buf, sizeof(buf), "BVH has %d or more inner levels: deeper than the %d-entry traversal stack (cudaUtils.h:22)", GPT_STACK_SIZE, GPT_STACK_SIZE);
        else snprintf(buf, sizeof(buf), "malformed BVH arrays (code %d): not a CudaBVH BVHLayout_Compact tree", r.bad_layout);
        return fail(c, GPT_ERR_LAYOUT, buf);
    }
    c->nodes = r.nodes; c->tris = r.tris; c->n_nodes = r.n_nodes; c->n_tris = r.n_tris; c->tree_depth = r.depth;
    // the same triangle references under a tree of this library's own builder (see k_env_rebuild); failure just leaves the one above
    const bool wanted = !c->own_tree && k_env_rebuild != 0 && !(c->flags & (GPT_FLAG_KEEP_TREE | GPT_FLAG_BINARY_BVH | GPT_FLAG_WIDE_BVH | GPT_FLAG_MEGAKERNEL)) &&
                        ((k_env_rebuild > 0 || k_env_split > 0) ? c->n_tris >= 2 : (c->n_tris >= 16 && c->n_tris <= (4 << 20)));
    if (wanted) {
        static const int k_env_ploc = []{ const char* v = getenv("GPT_PLOC_RADIUS"); return (v && *v) ? atoi(v) : 24; }();
        RepackResult a; int status = 0; float ms = 0.0f;
        static const int k_env_leaf = []{ const char* v = getenv("GPT_REBUILD_MAX_LEAF"); return (v && *v) ? atoi(v) : 8; }();
        e = gptk_rebuild(c->tris, c->n_tris, (const float4*)tris, k_env_leaf, (c->flags & GPT_FLAG_BUILD_LBVH) ? 0 : k_env_ploc, &a, &status, &ms, &c->launches, 0);
        if (e != cudaSuccess) cudaGetLastError();
        else if (status == 0) { c->alt_nodes = a.nodes; c->alt_tris = a.tris; c->alt_n_nodes = a.n_nodes; c->alt_n_tris = a.n_tris; c->alt_depth = a.depth; c->alt_build_ms = ms; }
        if (k_env_split != 0) {
            static const int k_env_split_leaf = []{ const char* v = getenv("GPT_SPLIT_MAX_LEAF"); return (v && *v) ? atoi(v) : 2; }();
            RepackResult sres; int sst = 0;
            e = gptk_split_leaves(c->nodes, c->n_nodes, c->tris, c->n_tris, (const float4*)tris, k_env_split_leaf, &sres, &sst, &c->launches, 0);
            if (e != cudaSuccess) cudaGetLastError();
            else if (sst == 0) { c->spl_nodes = sres.nodes; c->spl_tris = sres.tris; c->spl_n_nodes = sres.n_nodes; c->spl_depth = sres.depth; }
        }
    }
    return GPT_OK;
}

int gpt_upload_scene(gpt_ctx* c, const void* nodes, size_t n_node_vec4, const void* tris, size_t n_tri_vec4,
                     const int32_t* tri_indices, size_t n_indices, const gpt_sphere* spheres, size_t n_spheres,
                     const gpt_box* boxes, size_t n_boxes, int on_device)
{
    if (!c) return GPT_ERR_INVALID;
    if ((n_spheres && !spheres) || (n_boxes && !boxes)) return fail(c, GPT_ERR_INVALID, "gpt_upload_scene: NULL primitive array with non-zero count");
    if (n_spheres > (1u << 20) || n_boxes > (1u << 20)) return fail(c, GPT_ERR_INVALID, "gpt_upload_scene: too many primitives");
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaDeviceSynchronize());
    free_scene(c);
    // spheres / boxes may live on either side regardless of `on_device` (which describes the BVH arrays): UVA sorts it out
    const cudaMemcpyKind kind = cudaMemcpyDefault;
    int rc = GPT_OK;
    if (n_node_vec4 && nodes) {
        if (on_device) {
            rc = pack_bvh_from_device(c, nodes, n_node_vec4, tris, n_tri_vec4, tri_indices, n_indices);
        } else {
            if (!tris || !tri_indices) return fail(c, GPT_ERR_INVALID, "BVH nodes given without triangle / index arrays");
            void *dn = nullptr, *dt = nullptr; int32_t* di = nullptr;
            cudaError_t e;
            if ((e = cudaMalloc(&dn, n_node_vec4 * 16)) != cudaSuccess || (e = cudaMalloc(&dt, (n_tri_vec4 ? n_tri_vec4 : 1) * 16)) != cudaSuccess ||
                (e = cudaMalloc((void**)&di, (n_indices ? n_indices : 1) * 4)) != cudaSuccess ||
                (e = cudaMemcpy(dn, nodes, n_node_vec4 * 16, cudaMemcpyHostToDevice)) != cudaSuccess ||
                (e = cudaMemcpy(dt, tris, n_tri_vec4 * 16, cudaMemcpyHostToDevice)) != cudaSuccess ||
                (e = cudaMemcpy(di, tri_indices, n_indices * 4, cudaMemcpyHostToDevice)) != cudaSuccess) {
                cudaFree(dn); cudaFree(dt); cudaFree(di);
                return fail(c, GPT_ERR_CUDA, "gpt_upload_scene: staging copy", e);
            }
            rc = pack_bvh_from_device(c, dn, n_node_vec4, dt, n_tri_vec4, di, n_indices);
            cudaFree(dn); cudaFree(dt); cudaFree(di);
        }
        if (rc != GPT_OK) return rc;
    }
    if (n_spheres) {
        CU(c, cudaMalloc(&c->spheres, n_spheres * sizeof(gpt_sphere)));
        CU(c, cudaMemcpy(c->spheres, spheres, n_spheres * sizeof(gpt_sphere), kind));
        c->n_spheres = (int)n_spheres;
    }
    if (n_boxes) {
        CU(c, cudaMalloc(&c->boxes, n_boxes * sizeof(gpt_box)));
        CU(c, cudaMemcpy(c->boxes, boxes, n_boxes * sizeof(gpt_box), kind));
        c->n_boxes = (int)n_boxes;
    }
    c->has_scene = true;
    return GPT_OK;
}

int gpt_build_bvh(gpt_ctx* c, const float* verts, size_t n_verts, const int32_t* tris, size_t n_tris, int on_device,
                  uint32_t max_leaf_size, gpt_built_bvh* out)
{
    if (!c || !out) return GPT_ERR_INVALID;
    memset(out, 0, sizeof(*out));
    if (n_tris == 0) return fail(c, GPT_ERR_INVALID, "gpt_build_bvh: empty mesh");
    if (!verts || !tris || n_verts == 0) return fail(c, GPT_ERR_INVALID, "gpt_build_bvh: NULL vertex / triangle array");
    CU(c, cudaSetDevice(c->device));
    cudaEvent_t e0, e1;
    CU(c, cudaEventCreate(&e0)); CU(c, cudaEventCreate(&e1));
    cudaEventRecord(e0, 0);
    float* dv = nullptr; int* dt = nullptr;
    cudaError_t e = cudaSuccess;
    if (!on_device) {
        if ((e = cudaMalloc(&dv, n_verts * 12)) != cudaSuccess || (e = cudaMalloc(&dt, n_tris * 12)) != cudaSuccess ||
            (e = cudaMemcpy(dv, verts, n_verts * 12, cudaMemcpyHostToDevice)) != cudaSuccess ||
            (e = cudaMemcpy(dt, tris, n_tris * 12, cudaMemcpyHostToDevice)) != cudaSuccess) {
            cudaFree(dv); cudaFree(dt); cudaEventDestroy(e0); cudaEventDestroy(e1);
            return fail(c, GPT_ERR_CUDA, "gpt_build_bvh: staging copy", e);
        }
    }
    // topology: PLOC (area-driven clustering of the Morton-sorted triangles) unless GPT_FLAG_BUILD_LBVH / GPT_PLOC_RADIUS=0 ask for
    // the plain Morton radix tree; a PLOC tree deeper than the traversal stack (adversarial input) is rebuilt as a radix tree,
    // whose depth is bounded by the 63 key bits
    static const int k_env_ploc = []{ const char* v = getenv("GPT_PLOC_RADIUS"); return (v && *v) ? atoi(v) : 24; }();
    int radius = (c->flags & GPT_FLAG_BUILD_LBVH) ? 0 : k_env_ploc;
    BuildResult r;
    for (;;) {
        e = gptk_build_lbvh(on_device ? verts : dv, n_verts, on_device ? tris : dt, n_tris, max_leaf_size ? (int)max_leaf_size : 8, radius, &r, 0);
        c->launches += (unsigned long long)r.launches;
        if (e != cudaSuccess || radius <= 0) break;
        const bool deep = r.status == 0 && r.depth >= GPT_STACK_SIZE;      // sp reaches `depth`: int stack[GPT_STACK_SIZE] holds depth <= 63
        if (!deep && r.status != 4 && r.status != 5) break;
        if (deep) { cudaFree(r.nodes); cudaFree(r.tris); cudaFree(r.indices); }
        radius = 0;
    }
    cudaEventRecord(e1, 0);
    cudaEventSynchronize(e1);
    float ms = 0.0f; cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(dv); cudaFree(dt);
    if (e != cudaSuccess) return fail(c, GPT_ERR_CUDA, "gpt_build_bvh", e);
    if (r.status == 0 && r.depth >= GPT_STACK_SIZE) { cudaFree(r.nodes); cudaFree(r.tris); cudaFree(r.indices); r.status = 4; }
    if (r.status) {
        static const char* why[] = {"", "vertex index out of range", "non-finite vertex", "mesh too large for 32-bit offsets", "tree deeper than the 64-entry traversal stack (63 inner levels)", "clustering did not converge"};
        char buf[128]; snprintf(buf, sizeof(buf), "gpt_build_bvh: %s", why[r.status < 6 ? r.status : 0]);
        return fail(c, r.status == 3 || r.status == 4 ? GPT_ERR_LAYOUT : GPT_ERR_INVALID, buf);
    }
    c->built_nodes[c->built_next] = r.nodes; c->built_next = (c->built_next + 1) % 8;      // uploading these arrays needs no rebuild
    out->bvh.dev_triNode = r.nodes; out->bvh.triNodeSize = r.n_node_vec4;
    out->bvh.dev_triPtr = r.tris; out->bvh.triSize = r.n_tri_vec4;
    out->bvh.dev_triIndicesPtr = r.indices; out->bvh.triIndicesSize = r.n_indices;
    out->n_inner_nodes = (uint32_t)(r.n_node_vec4 / 4); out->n_leaves = (uint32_t)r.n_leaves;
    out->n_triangles = (uint32_t)n_tris; out->depth = (uint32_t)r.depth; out->build_ms = ms;
    return GPT_OK;
}

int gpt_free_bvh(gpt_ctx* c, gpt_built_bvh* b)
{
    if (!c || !b) return GPT_ERR_INVALID;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaDeviceSynchronize());
    if (c->src_nodes == b->bvh.dev_triNode) c->src_nodes = nullptr;      // the packed copy stays valid; force a re-pack if the address is reused
    cudaFree(b->bvh.dev_triNode); cudaFree(b->bvh.dev_triPtr); cudaFree(b->bvh.dev_triIndicesPtr);
    memset(b, 0, sizeof(*b));
    return GPT_OK;
}

int gpt_copy_to_host(gpt_ctx* c, void* dst, const void* src, size_t bytes)
{
    if (!c || (bytes && (!dst || !src))) return GPT_ERR_INVALID;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost));
    return GPT_OK;
}

int gpt_set_triangle_materials(gpt_ctx* c, const gpt_tri_material* mats, size_t n_mats, const int32_t* face_material, size_t n_faces)
{
    if (!c) return GPT_ERR_INVALID;
    if ((n_faces && (!mats || !face_material || n_mats == 0)) || n_mats > 0x7fffffffull || n_faces > 0x7fffffffull)
        return fail(c, GPT_ERR_INVALID, "gpt_set_triangle_materials: bad arguments");
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaDeviceSynchronize());
    cudaFree(c->tri_mats); cudaFree(c->face_mat);
    c->tri_mats = nullptr; c->face_mat = nullptr; c->n_tri_mats = 0; c->n_faces = 0;
    if (n_faces == 0) return GPT_OK;
    CU(c, cudaMalloc(&c->tri_mats, n_mats * sizeof(gpt_tri_material)));
    CU(c, cudaMalloc(&c->face_mat, n_faces * sizeof(int32_t)));
    CU(c, cudaMemcpy(c->tri_mats, mats, n_mats * sizeof(gpt_tri_material), cudaMemcpyDefault));
    CU(c, cudaMemcpy(c->face_mat, face_material, n_faces * sizeof(int32_t), cudaMemcpyDefault));
    c->n_tri_mats = (int)n_mats; c->n_faces = (int)n_faces;
    return GPT_OK;
}

static int make_params(gpt_ctx* c, const gpt_kernel_info* info, uint64_t seed, RenderParams* p)
{
    if (info->width < 2 || info->height < 2 || (long long)info->width * info->height > 0x7fffffffLL / 4)
        return fail(c, GPT_ERR_INVALID, "width/height out of range (need >= 2: the camera divides by w-1, h-1)");
    if (info->depth > 4096) return fail(c, GPT_ERR_INVALID, "depth out of range");
    p->width = info->width; p->height = info->height;
    for (int i = 0; i < 3; ++i) {
        p->front[i] = info->cam.front[i]; p->right[i] = info->cam.right[i]; p->up[i] = info->cam.up[i]; p->pos[i] = info->cam.pos[i];
        p->bk[i] = info->bkColor[i]; p->emi[i] = info->emi[i]; p->col[i] = info->col[i];
    }
    p->dist = info->cam.dist; p->aspect = info->cam.aspect; p->fov = info->cam.fov;
    p->cull = info->cullBackFaces ? 1 : 0;
    p->depth = info->depth;
    p->air = info->air_ref_index; p->glass = info->glass_ref_index;
    p->tri_mat = info->triCurrentMat;
    p->phong = info->phongExpo;
    p->seed = seed;
    p->flags = c->flags & (GPT_FLAG_NO_JITTER | GPT_FLAG_METAL_FIXED | GPT_FLAG_COUNT_BOUNCES);
    if (c->flags & GPT_FLAG_FAST) p->flags |= RP_FAST;
    return GPT_OK;
}

// builds the wide tree when it is forced or a candidate; a tree that cannot be quantised (non-finite boxes) or is too deep for
// the wide stack keeps the binary traversal -- same results either way
static void ensure_wide(gpt_ctx* c)
{
    if (c->wnodes || c->wide_tried || tree_choice(c) == 0) return;
    c->wide_tried = true;
    WideResult w;
    if (cudaDeviceSynchronize() != cudaSuccess) return;
    if (gptk_build_wide(c->nodes, c->n_nodes, &w, 0) != cudaSuccess) { cudaGetLastError(); return; }
    c->launches += (unsigned long long)w.launches;
    if (w.status != 0 || !w.wnodes || 3 * w.levels + 2 > GPT_W4_STACK) { cudaFree(w.wnodes); return; }
    c->wnodes = w.wnodes; c->n_wide = w.n_wide; c->wide_levels = w.levels;
}

static void uploaded_scene(gpt_ctx* c, DevScene* sc)
{
    ensure_wide(c);
    set_tree(c, sc, current_tree(c));                            // render calls refine this per call (render_wavefront)
    sc->spheres = c->spheres; sc->n_spheres = c->n_spheres;
    sc->boxes = c->boxes; sc->n_boxes = c->n_boxes;
    sc->tri_mats = c->tri_mats; sc->face_mat = c->face_mat; sc->n_tri_mats = c->n_tri_mats; sc->n_faces = c->n_faces;
}

int gpt_reorder_info(gpt_ctx* c, int* state, double* ns_per_ray_path_order, double* ns_per_ray_reordered)
{
    if (!c) return GPT_ERR_INVALID;
    const TraversalPlan pl = traversal_plan(c, (size_t)1 << 30);
    if (state) *state = (pl.trial && sort_choice(c) < 0) ? -1 : (current_sort(c) ? 1 : 0);
    if (ns_per_ray_path_order) *ns_per_ray_path_order = tree_ns(c, c->tune_done ? c->tuned_tree : current_tree(c));
    if (ns_per_ray_reordered) *ns_per_ray_reordered = c->tune_ns[2];
    return GPT_OK;
}

int gpt_rebuild_info(gpt_ctx* c, int* state, int* n_inner_nodes, int* depth, double* ns_per_ray, float* build_ms)
{
    if (!c) return GPT_ERR_INVALID;
    const int a = alt_choice(c);
    if (state) *state = a == 0 ? 0 : ((a < 0 && traversal_plan(c, (size_t)1 << 30).trial) ? -1 : (current_tree(c) == 2 ? 1 : 0));
    if (n_inner_nodes) *n_inner_nodes = c->alt_n_nodes;
    if (depth) *depth = c->alt_depth;
    if (ns_per_ray) *ns_per_ray = c->tune_ns[3];
    if (build_ms) *build_ms = c->alt_build_ms;
    return GPT_OK;
}

int gpt_tree_info(gpt_ctx* c, int* tree_in_use, int* pending, double ns_per_ray[4], int n_inner_nodes[4])
{
    if (!c) return GPT_ERR_INVALID;
    if (tree_in_use) *tree_in_use = current_tree(c);
    if (pending) *pending = traversal_plan(c, (size_t)1 << 30).trial ? 1 : 0;
    if (ns_per_ray) for (int i = 0; i < 4; ++i) ns_per_ray[i] = tree_ns(c, i);
    if (n_inner_nodes) { n_inner_nodes[0] = c->n_nodes; n_inner_nodes[1] = c->n_wide; n_inner_nodes[2] = c->alt_n_nodes; n_inner_nodes[3] = c->spl_n_nodes; }
    return GPT_OK;
}

int gpt_megakernel_tree_info(gpt_ctx* c, int* tree_in_use, int* pending, double ns_per_ray[4])
{
    if (!c) return GPT_ERR_INVALID;
    int ids[4];
    const int n = mega_cands(c, ids);
    if (tree_in_use) *tree_in_use = n == 1 ? ids[0] : (c->mega_done ? c->mega_tree : ((c->tune_done && c->tuned_tree != 1) ? c->tuned_tree : ids[0]));
    if (pending) *pending = (n > 1 && !c->mega_done && !(c->flags & GPT_FLAG_NO_TRIAL) && !k_env_no_trial) ? 1 : 0;
    if (ns_per_ray) for (int i = 0; i < 4; ++i) ns_per_ray[i] = c->mega_ns[i];
    return GPT_OK;
}

int gpt_trial_info(gpt_ctx* c, int* pending, double ns_per_ray[3])
{
    if (!c) return GPT_ERR_INVALID;
    if (pending) *pending = traversal_plan(c, (size_t)1 << 30).trial ? 1 : 0;
    if (ns_per_ray) { ns_per_ray[0] = c->tune_ns[0]; ns_per_ray[1] = c->tune_ns[1]; ns_per_ray[2] = c->tune_ns[2]; }
    return GPT_OK;
}

int gpt_wide_info(gpt_ctx* c, int* n_wide, int* levels)
{
    if (!c) return GPT_ERR_INVALID;
    CU(c, cudaSetDevice(c->device));
    ensure_wide(c);
    const bool on = current_wide(c) && c->wnodes;
    if (n_wide) *n_wide = on ? c->n_wide : 0;
    if (levels) *levels = on ? c->wide_levels : 0;
    return GPT_OK;
}

// scene for a launch: pointers carried by kernelInfo win (drop-in use), else the uploaded scene
static int launch_scene(gpt_ctx* c, const gpt_kernel_info* info, DevScene* sc)
{
    const gpt_bvh_data& b = info->bvhData;
    if (b.dev_triNode && b.triNodeSize) {
        const bool same = c->src_nodes == b.dev_triNode && c->src_tris == b.dev_triPtr && c->src_idx == b.dev_triIndicesPtr &&
                          c->src_n_nodes == b.triNodeSize && c->src_n_tris == b.triSize && c->src_n_idx == b.triIndicesSize;
        if (!same) {
            CU(c, cudaDeviceSynchronize());
            int rc = pack_bvh_from_device(c, b.dev_triNode, b.triNodeSize, b.dev_triPtr, b.triSize, b.dev_triIndicesPtr, b.triIndicesSize);
            if (rc != GPT_OK) return rc;
            c->src_nodes = b.dev_triNode; c->src_tris = b.dev_triPtr; c->src_idx = b.dev_triIndicesPtr;
            c->src_n_nodes = b.triNodeSize; c->src_n_tris = b.triSize; c->src_n_idx = b.triIndicesSize;
        }
    } else if (!c->has_scene && !(info->sphereTex && info->numSpheres)) {
        return fail(c, GPT_ERR_NO_SCENE, "no scene: call gpt_upload_scene or pass BVH/sphere device pointers in kernelInfo");
    }
    uploaded_scene(c, sc);
    if (info->sphereTex && info->numSpheres) {        // read live from the caller's array like the reference (tracer.cu:39)
        if (info->numSpheres > (1u << 20)) return fail(c, GPT_ERR_INVALID, "numSpheres out of range");
        sc->spheres = reinterpret_cast<const float*>(info->sphereTex); sc->n_spheres = (int)info->numSpheres;
    }
    return GPT_OK;
}

int gpt_invalidate_scene(gpt_ctx* c)
{
    if (!c) return GPT_ERR_INVALID;
    c->src_nodes = c->src_tris = c->src_idx = nullptr;      // the next gpt_launch that carries BVH pointers re-packs them
    return GPT_OK;
}

int gpt_launch(gpt_ctx* c, const gpt_kernel_info* info, gpt_stream stream)
{
    if (!c || !info) return GPT_ERR_INVALID;
    if (!info->accumBuffer) return fail(c, GPT_ERR_INVALID, "gpt_launch: accumBuffer is NULL");
    if (info->constantPdf == 0) return fail(c, GPT_ERR_INVALID, "gpt_launch: constantPdf must be >= 1");
    CU(c, cudaSetDevice(c->device));
    RenderParams p; DevScene sc;
    int rc = make_params(c, info, info->hash, &p); if (rc) return rc;
    rc = launch_scene(c, info, &sc); if (rc) return rc;
    if (!use_megakernel(c, (size_t)p.width * p.height)) {
        return render_wavefront(c, sc, p, nullptr, info->accumBuffer, info->dev_drawRes, info->constantPdf, 0ull, 1u, (cudaStream_t)stream);
    }
    return render_mega(c, sc, p, nullptr, info->accumBuffer, info->dev_drawRes, info->constantPdf, 0ull, 1u, (cudaStream_t)stream);
}

int gpt_render(gpt_ctx* c, const gpt_kernel_info* info, float* accum_f4, uint64_t seed, uint64_t sample_begin,
               uint32_t sample_count, gpt_stream stream)
{
    if (!c || !info || !accum_f4) return GPT_ERR_INVALID;
    if (sample_count == 0) return GPT_OK;
    CU(c, cudaSetDevice(c->device));
    RenderParams p; DevScene sc;
    int rc = make_params(c, info, seed, &p); if (rc) return rc;
    rc = launch_scene(c, info, &sc); if (rc) return rc;
    if (!use_megakernel(c, (size_t)p.width * p.height * sample_count)) {
        return render_wavefront(c, sc, p, reinterpret_cast<float4*>(accum_f4), nullptr, nullptr, 1ull, sample_begin, sample_count, (cudaStream_t)stream);
    }
    return render_mega(c, sc, p, reinterpret_cast<float4*>(accum_f4), nullptr, nullptr, 1ull, sample_begin, sample_count, (cudaStream_t)stream);
}

int gpt_resolve(gpt_ctx* c, const float* accum_f4, int width, int height, uint32_t* out_rgba8, float* out_rgb_f32, gpt_stream stream)
{
    if (!c || !accum_f4 || width <= 0 || height <= 0) return GPT_ERR_INVALID;
    CU(c, cudaSetDevice(c->device));
    KernelTimer* kt = (c->flags & GPT_FLAG_TIME_KERNELS) ? &c->timer : nullptr;
    if (kt) kt->tick(KT_RESOLVE, (cudaStream_t)stream);
    cudaError_t e = gptk_resolve(reinterpret_cast<const float4*>(accum_f4), width * height, out_rgba8, out_rgb_f32, c->sm_count, (cudaStream_t)stream);
    if (kt) kt->tick(KT_END, (cudaStream_t)stream);
    if (e != cudaSuccess) return fail(c, GPT_ERR_CUDA, "resolve launch", e);
    c->launches += 1;
    return GPT_OK;
}

int gpt_render_frame_host(gpt_ctx* c, const gpt_kernel_info* info, uint64_t seed, uint64_t sample_begin, uint32_t spp,
                          uint32_t* host_rgba8, float* host_accum_f4)
{
    if (!c || !info) return GPT_ERR_INVALID;
    CU(c, cudaSetDevice(c->device));
    const size_t n = (size_t)info->width * (size_t)info->height;
    if (n == 0) return fail(c, GPT_ERR_INVALID, "empty frame");
    if (c->frame_pixels != n) {
        cudaFree(c->frame_accum); cudaFree(c->frame_rgba); c->frame_accum = nullptr; c->frame_rgba = nullptr; c->frame_pixels = 0;
        CU(c, cudaMalloc(&c->frame_accum, n * sizeof(float4)));
        CU(c, cudaMalloc(&c->frame_rgba, n * sizeof(uint32_t)));
        c->frame_pixels = n;
    }
    CU(c, cudaMemsetAsync(c->frame_accum, 0, n * sizeof(float4), 0));
    int rc = gpt_render(c, info, reinterpret_cast<float*>(c->frame_accum), seed, sample_begin, spp, 0); if (rc) return rc;
    rc = gpt_resolve(c, reinterpret_cast<const float*>(c->frame_accum), info->width, info->height, c->frame_rgba, nullptr, 0); if (rc) return rc;
    if (host_rgba8) CU(c, cudaMemcpyAsync(host_rgba8, c->frame_rgba, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, 0));
    if (host_accum_f4) CU(c, cudaMemcpyAsync(host_accum_f4, c->frame_accum, n * sizeof(float4), cudaMemcpyDeviceToHost, 0));
    CU(c, cudaStreamSynchronize(0));
    return GPT_OK;
}

int gpt_trace_primary(gpt_ctx* c, const gpt_kernel_info* info, uint64_t seed, uint64_t sample,
                      int32_t* geom, int32_t* tri_idx, int32_t* prim_idx, float* t, gpt_stream stream)
{
    if (!c || !info) return GPT_ERR_INVALID;
    CU(c, cudaSetDevice(c->device));
    RenderParams p; DevScene sc;
    int rc = make_params(c, info, seed, &p); if (rc) return rc;
    rc = launch_scene(c, info, &sc); if (rc) return rc;
    cudaError_t e = gptk_primary(sc, p, sample, geom, tri_idx, prim_idx, t, c->sm_count, (cudaStream_t)stream);
    if (e != cudaSuccess) return fail(c, GPT_ERR_CUDA, "primary launch", e);
    c->launches += 1;
    return GPT_OK;
}

int gpt_trace_rays(gpt_ctx* c, const float* o, const float* d, size_t n, int cull, int32_t* tri_idx, int32_t* tri_addr,
                   float* t, gpt_stream stream)
{
    if (!c || (n && (!o || !d))) return GPT_ERR_INVALID;
    if (!c->has_scene) return fail(c, GPT_ERR_NO_SCENE, "gpt_trace_rays: no uploaded scene");
    CU(c, cudaSetDevice(c->device));
    DevScene sc; uploaded_scene(c, &sc);
    cudaError_t e = gptk_trace_rays(sc, o, d, n, cull, tri_idx, tri_addr, t, c->sm_count, (cudaStream_t)stream);
    if (e != cudaSuccess) return fail(c, GPT_ERR_CUDA, "trace_rays launch", e);
    if (n) c->launches += 1;
    return GPT_OK;
}

int gpt_get_stats(gpt_ctx* c, gpt_stats* out)
{
    if (!c || !out) return GPT_ERR_INVALID;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaDeviceSynchronize());
    unsigned long long h[32];
    CU(c, cudaMemcpy(h, c->stats, sizeof(h), cudaMemcpyDeviceToHost));
    out->rays = h[0]; out->paths = h[1];
    for (int i = 0; i < 16; ++i) out->rays_by_bounce[i] = h[2 + i];
    out->kernel_launches = c->launches;
    return GPT_OK;
}

int gpt_frame_free(gpt_ctx* c)
{
    if (!c) return GPT_ERR_INVALID;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    for (int r = 0; r < 16; ++r) { if (c->peer_base[r]) cudaIpcCloseMemHandle(c->peer_base[r]); c->peer_base[r] = nullptr; }
    c->peer_world = 0;
    cudaFree(c->frames); c->frames = nullptr; c->frame_w = c->frame_h = c->frame_n = 0;
    return GPT_OK;
}

int gpt_frame_alloc(gpt_ctx* c, int width, int height, int n_buffers, gpt_ipc_mem_handle* handle_out)
{
    if (!c || width <= 0 || height <= 0 || n_buffers < 1 || n_buffers > 8) return GPT_ERR_INVALID;
    gpt_frame_free(c);
    CU(c, cudaSetDevice(c->device));
    const size_t bytes = sizeof(float4) * (size_t)width * height * n_buffers;
    CU(c, cudaMalloc(&c->frames, bytes));
    CU(c, cudaMemset(c->frames, 0, bytes));
    c->frame_w = width; c->frame_h = height; c->frame_n = n_buffers;
    if (handle_out) {
        static_assert(sizeof(cudaIpcMemHandle_t) == sizeof(gpt_ipc_mem_handle), "IPC handle size");
        CU(c, cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(handle_out), c->frames));
    }
    return GPT_OK;
}

int gpt_frame_open_peers(gpt_ctx* c, const gpt_ipc_mem_handle* handles, int world, int rank)
{
    if (!c || !handles || world < 1 || world > 16 || rank < 0 || rank >= world) return GPT_ERR_INVALID;
    if (!c->frames) return fail(c, GPT_ERR_INVALID, "gpt_frame_open_peers: call gpt_frame_alloc first");
    CU(c, cudaSetDevice(c->device));
    for (int r = 0; r < world; ++r) {
        if (r == rank) continue;
        cudaIpcMemHandle_t h; memcpy(&h, handles + r, sizeof(h));
        CU(c, cudaIpcOpenMemHandle(&c->peer_base[r], h, cudaIpcMemLazyEnablePeerAccess));
    }
    c->peer_world = world; c->peer_rank = rank;
    return GPT_OK;
}

float* gpt_frame_ptr(gpt_ctx* c, int index)
{
    if (!c || !c->frames || index < 0 || index >= c->frame_n) return nullptr;
    return reinterpret_cast<float*>(c->frames + (size_t)index * c->frame_w * c->frame_h);
}

int gpt_frame_clear(gpt_ctx* c, int index, gpt_stream stream)
{
    float* p = gpt_frame_ptr(c, index);
    if (!p) return GPT_ERR_INVALID;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaMemsetAsync(p, 0, sizeof(float4) * (size_t)c->frame_w * c->frame_h, (cudaStream_t)stream));
    return GPT_OK;
}

int gpt_reduce_resolve_ptrs(gpt_ctx* c, const float* const* frames, int n_frames, int width, int height,
                            float* total, uint32_t* out_rgba8, float* out_rgb, gpt_stream stream)
{
    if (!c || !frames || n_frames < 1 || n_frames > 16 || width <= 0 || height <= 0) return GPT_ERR_INVALID;
    for (int r = 0; r < n_frames; ++r) if (!frames[r]) return GPT_ERR_INVALID;
    CU(c, cudaSetDevice(c->device));
    KernelTimer* kt = (c->flags & GPT_FLAG_TIME_KERNELS) ? &c->timer : nullptr;
    if (kt) kt->tick(KT_RESOLVE, (cudaStream_t)stream);
    cudaError_t e = gptk_reduce_resolve(reinterpret_cast<const float4* const*>(frames), n_frames, width * height, reinterpret_cast<float4*>(total),
                                        out_rgba8, out_rgb, c->sm_count, (cudaStream_t)stream);
    if (kt) kt->tick(KT_END, (cudaStream_t)stream);
    if (e != cudaSuccess) return fail(c, GPT_ERR_CUDA, "reduce_resolve launch", e);
    c->launches += 1;
    return GPT_OK;
}

int gpt_reduce_resolve(gpt_ctx* c, int index, float* total, uint32_t* out_rgba8, float* out_rgb, gpt_stream stream)
{
    if (!c || !c->frames || index < 0 || index >= c->frame_n) return GPT_ERR_INVALID;
    const int world = c->peer_world ? c->peer_world : 1;
    const size_t off = (size_t)index * c->frame_w * c->frame_h;
    const float* ptrs[16];
    for (int r = 0; r < world; ++r) {
        const float4* base = (world == 1 || r == c->peer_rank) ? c->frames : reinterpret_cast<const float4*>(c->peer_base[r]);
        if (!base) return fail(c, GPT_ERR_INVALID, "gpt_reduce_resolve: peer not opened");
        ptrs[r] = reinterpret_cast<const float*>(base + off);
    }
    return gpt_reduce_resolve_ptrs(c, ptrs, world, c->frame_w, c->frame_h, total, out_rgba8, out_rgb, stream);
}

int gpt_get_kernel_times(gpt_ctx* c, gpt_kernel_times* out)
{
    if (!c || !out) return GPT_ERR_INVALID;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaDeviceSynchronize());
    c->timer.collect(c->kt_ms, c->kt_n);
    for (int i = 0; i < 8; ++i) { out->ms[i] = c->kt_ms[i]; out->launches[i] = c->kt_n[i]; }
    return GPT_OK;
}

int gpt_measure_peaks(gpt_ctx* c, gpt_peaks* out)
{
    if (!c || !out) return GPT_ERR_INVALID;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaDeviceSynchronize());
    cudaError_t e = gptk_measure_peaks(c->sm_count, &out->gather_l1_gbs, &out->gather_l2_gbs, &out->gather_hbm_gbs, &out->warp_inst_per_s,
                                       &out->stream_l1_gbs, &out->stream_l2_gbs);
    if (e != cudaSuccess) return fail(c, GPT_ERR_CUDA, "gpt_measure_peaks", e);
    c->launches += 16;
    return GPT_OK;
}

int gpt_reset_stats(gpt_ctx* c)
{
    if (!c) return GPT_ERR_INVALID;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaDeviceSynchronize());
    CU(c, cudaMemset(c->stats, 0, 32 * sizeof(unsigned long long)));
    c->launches = 0;
    c->timer.used = 0;
    for (int i = 0; i < KT_CLASSES; ++i) { c->kt_ms[i] = 0.0; c->kt_n[i] = 0; }
    return GPT_OK;
}

} // extern "C"

// ---- gpt_render_static: the wavefront pipeline on scratch that is never allocated ----------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc (code under
// CUDA graph capture, tight serving loops, harnesses that forbid it).  The per-path state and per-pixel sums that ensure_wave()
// cudaMallocs for gpt_render live here in ONE file-scope device array of fixed size, so every GPU that loads the library holds
// GPT_STATIC_SCRATCH_BYTES (30 MiB) until the process exits.  The array's name is a host-side stand-in in host code, not a device
// address: static_scratch_base() asks the runtime for the address once per device (cudaGetSymbolAddress with that device current)
// and the kernels only ever see that pointer.
//   layout (offsets are multiples of 256 B; P = GPT_STATIC_MAX_PATHS paths per wave, X = GPT_STATIC_MAX_PIXELS pixels):
//     ray 2 x 32 P | accu 16 P | lsum 16 X | hit 8 P | mask_z 2 x 4 P | queue[0] 4 P | queue[1] 4 P | counters 4 KiB
// All calls on a device share the array: the queue counters are cleared on `stream` at the start of every call (everything else is
// written before it is read), and two calls that overlap on the GPU (different streams, or different contexts of one device) race.
#define GPT_STATIC_P ((size_t)GPT_STATIC_MAX_PATHS)
#define GPT_STATIC_X ((size_t)GPT_STATIC_MAX_PIXELS)
#define GPT_STATIC_OFF_RAY      ((size_t)0)
#define GPT_STATIC_OFF_ACCU     (GPT_STATIC_OFF_RAY + 64 * GPT_STATIC_P)
#define GPT_STATIC_OFF_LSUM     (GPT_STATIC_OFF_ACCU + 16 * GPT_STATIC_P)
#define GPT_STATIC_OFF_HIT      (GPT_STATIC_OFF_LSUM + 16 * GPT_STATIC_X)
#define GPT_STATIC_OFF_MASK     (GPT_STATIC_OFF_HIT + 8 * GPT_STATIC_P)
#define GPT_STATIC_OFF_Q0       (GPT_STATIC_OFF_MASK + 8 * GPT_STATIC_P)
#define GPT_STATIC_OFF_Q1       (GPT_STATIC_OFF_Q0 + 4 * GPT_STATIC_P)
#define GPT_STATIC_OFF_COUNTERS (GPT_STATIC_OFF_Q1 + 4 * GPT_STATIC_P)
#define GPT_STATIC_COUNTER_BYTES ((size_t)4096)
#define GPT_STATIC_SCRATCH_BYTES (GPT_STATIC_OFF_COUNTERS + GPT_STATIC_COUNTER_BYTES)
static_assert(GPT_STATIC_MAX_PATHS % 64 == 0 && GPT_STATIC_MAX_PIXELS % 16 == 0, "sub-array offsets must stay 256-byte aligned");
static_assert(2 * (GPT_STATIC_MAX_DEPTH + 1) * sizeof(unsigned int) <= GPT_STATIC_COUNTER_BYTES, "counter block too small for the depth limit");

static __device__ __align__(256) unsigned char g_static_scratch[GPT_STATIC_SCRATCH_BYTES];

#define GPT_STATIC_MAX_DEVICES 64
static unsigned char* g_static_base[GPT_STATIC_MAX_DEVICES];      // device address of g_static_scratch on each device, nullptr = not asked yet
static std::mutex g_static_mutex;

// device address of the array on `device`, which must be the current device; nullptr + *err on failure
static unsigned char* static_scratch_base(int device, cudaError_t* err)
{
    *err = cudaSuccess;
    if (device < 0 || device >= GPT_STATIC_MAX_DEVICES) { *err = cudaErrorInvalidDevice; return nullptr; }
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!g_static_base[device]) {
        void* ptr = nullptr;
        *err = cudaGetSymbolAddress(&ptr, g_static_scratch);
        if (*err != cudaSuccess || !ptr) { if (*err == cudaSuccess) *err = cudaErrorInvalidSymbol; return nullptr; }
        g_static_base[device] = static_cast<unsigned char*>(ptr);
    }
    return g_static_base[device];
}

extern "C" int gpt_render_static(gpt_ctx* c, const gpt_kernel_info* info, float* accum_f4, uint64_t seed, uint64_t sample_begin,
                                 uint32_t sample_count, gpt_stream stream)
{
    if (!c || !info || !accum_f4) return GPT_ERR_INVALID;
    if (sample_count == 0) return GPT_OK;
    CU(c, cudaSetDevice(c->device));
    char buf[256];
    const long long n_pix = (long long)info->width * (long long)info->height;
    if (info->width < 2 || info->height < 2 || n_pix > (long long)GPT_STATIC_MAX_PIXELS) {
        snprintf(buf, sizeof(buf), "gpt_render_static: frame of %d x %d pixels is outside the static scratch (2 x 2 .. %d pixels = GPT_STATIC_MAX_PIXELS); use gpt_render",
                 info->width, info->height, (int)GPT_STATIC_MAX_PIXELS);
        return fail(c, GPT_ERR_INVALID, buf);
    }
    if (info->depth > GPT_STATIC_MAX_DEPTH) {
        snprintf(buf, sizeof(buf), "gpt_render_static: depth %u is beyond GPT_STATIC_MAX_DEPTH = %d; use gpt_render", info->depth, (int)GPT_STATIC_MAX_DEPTH);
        return fail(c, GPT_ERR_INVALID, buf);
    }
    if (!c->has_scene) return fail(c, GPT_ERR_NO_SCENE, "gpt_render_static: no uploaded scene (gpt_upload_scene is the set-up call; BVH pointers in kernelInfo are not packed here, that allocates)");
    RenderParams p;
    int rc = make_params(c, info, seed, &p); if (rc) return rc;
    // the tree in use (uploaded / rebuilt at upload / 4-wide); the 4-wide tree only if an earlier call already built it (building it allocates), never a trial
    DevScene sc;
    { const int tree = current_tree(c); set_tree(c, &sc, (tree == 1 && !c->wnodes) ? 0 : tree); }
    sc.spheres = c->spheres; sc.n_spheres = c->n_spheres;
    sc.boxes = c->boxes; sc.n_boxes = c->n_boxes;
    sc.tri_mats = c->tri_mats; sc.face_mat = c->face_mat; sc.n_tri_mats = c->n_tri_mats; sc.n_faces = c->n_faces;
    if (info->sphereTex && info->numSpheres) {        // read live from the caller's array, as gpt_render does (tracer.cu:39)
        if (info->numSpheres > (1u << 20)) return fail(c, GPT_ERR_INVALID, "numSpheres out of range");
        sc.spheres = reinterpret_cast<const float*>(info->sphereTex); sc.n_spheres = (int)info->numSpheres;
    }
    cudaError_t e;
    unsigned char* base = static_scratch_base(c->device, &e);
    if (!base) return fail(c, GPT_ERR_CUDA, "gpt_render_static: cudaGetSymbolAddress(g_static_scratch)", e);
    WaveBuffers wb = {};
    wb.ray = reinterpret_cast<float4*>(base + GPT_STATIC_OFF_RAY);
    wb.accu = reinterpret_cast<float4*>(base + GPT_STATIC_OFF_ACCU);
    wb.lsum = reinterpret_cast<float4*>(base + GPT_STATIC_OFF_LSUM);
    wb.hit = reinterpret_cast<int2*>(base + GPT_STATIC_OFF_HIT);
    wb.mask_z = reinterpret_cast<float*>(base + GPT_STATIC_OFF_MASK);
    wb.queue[0] = reinterpret_cast<uint32_t*>(base + GPT_STATIC_OFF_Q0);
    wb.queue[1] = reinterpret_cast<uint32_t*>(base + GPT_STATIC_OFF_Q1);
    wb.counters = reinterpret_cast<unsigned int*>(base + GPT_STATIC_OFF_COUNTERS);
    wb.capacity = GPT_STATIC_P;                       // calls of more paths run as several waves (gptk_render_wavefront)
    if (!c->wave_cfg.ready) gptk_wave_launch_cfg(&c->wave_cfg, c->sm_count);
    cudaStream_t st = (cudaStream_t)stream;
    // the array keeps the last call's contents (possibly of a call that failed half-way): the counters must start at zero
    CU(c, cudaMemsetAsync(wb.counters, 0, gptk_wave_counter_count(p.depth) * sizeof(unsigned int), st));
    e = gptk_render_wavefront(sc, p, wb, c->wave_cfg, reinterpret_cast<float4*>(accum_f4), nullptr, nullptr, 1ull, sample_begin, sample_count,
                              c->stats, st, &c->launches, nullptr);
    if (e != cudaSuccess) return fail(c, GPT_ERR_CUDA, "gpt_render_static: launch", e);
    return GPT_OK;
}
