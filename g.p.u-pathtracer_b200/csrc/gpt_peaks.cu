// gpt_peaks.cu -- micro-benchmarks that measure, on the device the context runs on, the peaks the traversal kernels are
// actually bounded by (gpt_measure_peaks; bench.py states its roofline fractions against THESE, not against data-sheet figures):
//   gather bandwidth  random 64-byte fetches (ldg_node64: 2 x LDG.256 through the read-only path, the BVH node fetch of
//                     gpt_traverse.cuh node_next) from a working set that fits L1 (64 KB), fits L2 (24 MB), or only HBM (2 GB);
//                     independent addresses, 128-thread CTAs at the extend kernel's grid (SMs x 10)
//   issue rate        eight independent FFMA chains per thread: warp instructions issued per second with every scheduler busy
//                     (= SMs x 4 schedulers x SM clock when nothing stalls)
// Diagnostics only: nothing on the render path calls this.
#include "gpt_kernels.h"

__global__ void __launch_bounds__(128)
k_peak_gather(const float4* __restrict__ buf, const unsigned mask, const int iters, float4* __restrict__ sink)
{
    unsigned idx = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (int i = 0; i < iters; ++i) {
        idx = idx * 1664525u + 1013904223u;
        const float4* p = buf + 4 * (size_t)((idx >> 7) & mask);
        float4 a, b, c, d;
        ldg_node64(p, a, b, c, d);                 // the node fetch of gpt_traverse.cuh, same load instructions
        acc.x += a.x + b.x; acc.y += a.y + c.y; acc.z += b.z + d.z; acc.w += c.w + d.w;
    }
    if (acc.x == 123.456f) sink[0] = acc;      // never true for the zeroed buffer: keeps the loads alive
}

// coalesced 16-byte loads sweeping a working set again and again: the hard ceiling of a cache level's bandwidth
__global__ void __launch_bounds__(256)
k_peak_stream(const float4* __restrict__ buf, const unsigned mask, const int iters, float4* __restrict__ sink)
{
    unsigned idx = blockIdx.x * 8191u * blockDim.x + threadIdx.x;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 8
    for (int i = 0; i < iters; ++i) {
        const float4 a = __ldg(buf + (idx & mask));
        acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w;
        idx += blockDim.x;
    }
    if (acc.x == 123.456f) sink[0] = acc;
}

__global__ void __launch_bounds__(128)
k_peak_issue(const int iters, const float seed, float* __restrict__ sink)
{
    float a0 = seed, a1 = seed + 1.f, a2 = seed + 2.f, a3 = seed + 3.f, a4 = seed + 4.f, a5 = seed + 5.f, a6 = seed + 6.f, a7 = seed + 7.f;
    const float m = 0.999f, c = 0.001f;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            a0 = fmaf(a0, m, c); a1 = fmaf(a1, m, c); a2 = fmaf(a2, m, c); a3 = fmaf(a3, m, c);
            a4 = fmaf(a4, m, c); a5 = fmaf(a5, m, c); a6 = fmaf(a6, m, c); a7 = fmaf(a7, m, c);
        }
    }
    const float s = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
    if (s == 123.456f) sink[0] = s;
}

static cudaError_t time_kernel(cudaEvent_t e0, cudaEvent_t e1, float* best_ms, int reps, void (*launch)(void*), void* arg)
{
    *best_ms = 1e30f;
    for (int r = 0; r < reps + 1; ++r) {
        cudaEventRecord(e0, 0);
        launch(arg);
        cudaEventRecord(e1, 0);
        cudaError_t e = cudaEventSynchronize(e1);
        if (e != cudaSuccess) return e;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        if (r > 0 && ms < *best_ms) *best_ms = ms;       // first run = warm-up (also warms the cache level under test)
    }
    return cudaGetLastError();
}

struct GatherArg { const float4* buf; unsigned mask; int iters; int grid; float4* sink; };
static void launch_gather(void* p) { GatherArg* a = (GatherArg*)p; k_peak_gather<<<a->grid, 128>>>(a->buf, a->mask, a->iters, a->sink); }
static void launch_stream(void* p) { GatherArg* a = (GatherArg*)p; k_peak_stream<<<a->grid, 256>>>(a->buf, a->mask, a->iters, a->sink); }
struct IssueArg { int iters; int grid; float* sink; };
static void launch_issue(void* p) { IssueArg* a = (IssueArg*)p; k_peak_issue<<<a->grid, 128>>>(a->iters, 1.0f, a->sink); }

cudaError_t gptk_measure_peaks(int sm_count, double* l1_gbs, double* l2_gbs, double* hbm_gbs, double* warp_inst_per_s,
                               double* stream_l1_gbs, double* stream_l2_gbs)
{
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    float4* buf = nullptr; float4* sink = nullptr;
    cudaError_t e;
    const size_t big = (size_t)2 << 30;
#define CK(x) do { e = (x); if (e != cudaSuccess) goto done; } while (0)
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    CK(cudaMalloc(&buf, big)); CK(cudaMalloc(&sink, 64));
    CK(cudaMemset(buf, 0, big));
    {
        const int grid = sm_count * 10;
        const double threads = (double)grid * 128.0;
        float ms;
        GatherArg g = {buf, (unsigned)((64u << 10) / 64 - 1), 4096, grid, sink};
        CK(time_kernel(e0, e1, &ms, 3, launch_gather, &g));
        *l1_gbs = threads * g.iters * 64.0 / (ms * 1e-3) / 1e9;
        g.mask = (unsigned)(((size_t)16 << 20) / 64 - 1); g.iters = 1024;      // 16 MiB: inside one L2 partition's share
        CK(time_kernel(e0, e1, &ms, 3, launch_gather, &g));
        *l2_gbs = threads * g.iters * 64.0 / (ms * 1e-3) / 1e9;
        g.mask = (unsigned)(big / 64 - 1); g.iters = 256;
        CK(time_kernel(e0, e1, &ms, 3, launch_gather, &g));
        *hbm_gbs = threads * g.iters * 64.0 / (ms * 1e-3) / 1e9;
        // streaming ceilings: every CTA sweeps its own 32 KiB window (L1), all CTAs sweep the same 16 MiB (L2; L1 holds 0.2 % of it)
        GatherArg s1 = {buf, (unsigned)((32u << 10) / 16 - 1), 8192, sm_count * 8, sink};
        CK(time_kernel(e0, e1, &ms, 3, launch_stream, &s1));
        *stream_l1_gbs = (double)s1.grid * 256.0 * s1.iters * 16.0 / (ms * 1e-3) / 1e9;
        GatherArg s2 = {buf, (unsigned)(((size_t)16 << 20) / 16 - 1), 2048, sm_count * 8, sink};
        CK(time_kernel(e0, e1, &ms, 3, launch_stream, &s2));
        *stream_l2_gbs = (double)s2.grid * 256.0 * s2.iters * 16.0 / (ms * 1e-3) / 1e9;
        IssueArg ia = {2048, sm_count * 16, reinterpret_cast<float*>(sink)};
        CK(time_kernel(e0, e1, &ms, 3, launch_issue, &ia));
        *warp_inst_per_s = ((double)ia.grid * 4.0) * ((double)ia.iters * 64.0) / (ms * 1e-3);     // 4 warps per CTA, 64 FFMA per trip
    }
done:
    cudaFree(buf); cudaFree(sink);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    return e;
#undef CK
}
