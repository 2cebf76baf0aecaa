// gpt_wide.cu -- device-side collapse of the packed binary BVH (gpt_repack.cu / gpt_build.cu) into a 4-wide BVH with
// 8-bit quantised child boxes: one 64-B node describes FOUR children where the binary layout describes two
// ("collapse to BVH4/8 + quantised/compressed nodes").  The triangle array is shared with the binary tree.
//
// Why: on large scenes the extend kernel is bound by L1 request throughput and by the number of dependent fetch rounds
// (ncu on 20 M triangles: L1 90 %, DRAM 2 %, 36 tree levels) -- not by HBM.  A 64-B wide node halves both.
//
// Node, 4 x float4 (64 B, one 64-B aligned fetch group like the binary node):
//   [0] = origin.x origin.y origin.z  bits(ex | ey<<8 | ez<<16)   e* = biased exponent: cell size on that axis = 2^(e-127)
//   [1] = u32 lo.x[4]  hi.x[4]  lo.y[4]  hi.y[4]                   one byte per child: plane = origin + q * cell (one cell of
//                                                                  slack outwards, see k_w4_collapse_level); unused slot: lo 255, hi 0
//   [2] = u32 lo.z[4]  hi.z[4]  0  0
//   [3] = int child[4]     >= 0: float4 index of a wide node;  < 0: ~(first packed triangle of a leaf);  GPT_W4_EMPTY: unused
// Quantisation is conservative (lo rounded down, hi up, verified in float arithmetic), so a quantised box CONTAINS the
// binary tree's box: the wide traversal visits a superset of the leaves the binary one visits and reports the same
// nearest hit.  What may differ: which of two bit-equal t's wins (visit order), i.e. which COPY of a spatially split
// triangle is reported -- never t or the original face index.
//
// Collapse: top-down, one launch per wide level.  A wide node starts from a binary node's two children and, while it has
// a free slot, replaces its inner child of largest surface area by that child's two children (the usual greedy
// surface-area collapse).  Levels are laid out one after the other (breadth first), children of a level contiguous.
#include "gpt_kernels.h"
#include <math.h>

struct W4Slot { float lo[3], hi[3]; int child; };

__device__ __forceinline__ void w4_read_pair(const float4* __restrict__ nodes2, const int f4, W4Slot& a, W4Slot& b)
{
    const float4 n0 = nodes2[f4], n1 = nodes2[f4 + 1], nz = nodes2[f4 + 2];
    const int4 ch = *reinterpret_cast<const int4*>(nodes2 + f4 + 3);
    a.lo[0] = n0.x; a.hi[0] = n0.y; a.lo[1] = n0.z; a.hi[1] = n0.w; a.lo[2] = nz.x; a.hi[2] = nz.y; a.child = ch.x;
    b.lo[0] = n1.x; b.hi[0] = n1.y; b.lo[1] = n1.z; b.hi[1] = n1.w; b.lo[2] = nz.z; b.hi[2] = nz.w; b.child = ch.y;
}

__device__ __forceinline__ float w4_area(const W4Slot& s)
{
    const float dx = s.hi[0] - s.lo[0], dy = s.hi[1] - s.lo[1], dz = s.hi[2] - s.lo[2];
    return dx * dy + dy * dz + dz * dx;
}

__global__ void __launch_bounds__(128)
k_w4_collapse_level(const float4* __restrict__ nodes2, const int* __restrict__ q_in, const int n_in, const int base_in,
                    const int base_next, int* __restrict__ q_out, int* __restrict__ n_out, float4* __restrict__ wnodes,
                    int* __restrict__ bad)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_in) return;
    W4Slot s[4];
    int cnt = 2;
    w4_read_pair(nodes2, q_in[i], s[0], s[1]);
    while (cnt < 4) {
        int best = -1; float best_area = -1.0f;
        for (int k = 0; k < cnt; ++k)
            if (s[k].child >= 0) { const float a = w4_area(s[k]); if (a > best_area) { best_area = a; best = k; } }
        if (best < 0) {      // no inner child left -- or only ones with a NaN area: take the first inner child, if any
            for (int k = 0; k < cnt; ++k) if (s[k].child >= 0) { best = k; break; }
            if (best < 0) break;
        }
        W4Slot a, b;
        w4_read_pair(nodes2, s[best].child, a, b);
        s[best] = a; s[cnt++] = b;
    }
    // frame of the node: per axis the smallest power-of-two cell with 253 cells covering the extent of the union box, origin one
    // cell BELOW the union's lower corner.  A child's planes are the tightest conservative ones (plane(q) = origin + q * cell,
    // verified in float arithmetic) moved ONE MORE cell outwards: the traversal evaluates (2^23 + q) * s + (b - 2^23 * s) with
    // the byte dropped into the mantissa of 2^23 -- no int-to-float conversion, no subtraction -- and rounding b - 2^23 * s
    // costs up to half a cell, which the extra cell absorbs.
    float org[3], cell[3];
    unsigned ebits = 0;
    bool ok = true;
    for (int ax = 0; ax < 3; ++ax) {
        float lo = s[0].lo[ax], hi = s[0].hi[ax];
        for (int k = 1; k < cnt; ++k) { lo = fminf(lo, s[k].lo[ax]); hi = fmaxf(hi, s[k].hi[ax]); }
        for (int k = 0; k < cnt; ++k) ok = ok && isfinite(s[k].lo[ax]) && isfinite(s[k].hi[ax]) && s[k].lo[ax] <= s[k].hi[ax];
        const float ext = hi - lo;
        int e = -100;
        if (ext > 0.0f && isfinite(ext)) { (void)frexpf(ext / 253.0f, &e); }
        else if (!(ext == 0.0f)) ok = false;
        // a cell must stay above the rounding of the coordinates themselves, or origin - cell == origin
        const float mag = fmaxf(fabsf(lo), fabsf(hi));
        if (mag > 0.0f && isfinite(mag)) { int em; (void)frexpf(mag, &em); if (e < em - 22) e = em - 22; }
        if (e < -100) e = -100;
        if (e > 100) { ok = false; e = 100; }
        float c = ldexpf(1.0f, e), o = lo - c;
        while (ok) {
            o = lo - c;
            while (fmaf(1.0f, c, o) > lo) o = nextafterf(o, -INFINITY);      // plane(1) must not lie above the lower corner
            if (fmaf(254.0f, c, o) >= hi && o < lo) break;
            if (++e > 100) { ok = false; break; }
            c = ldexpf(1.0f, e);
        }
        cell[ax] = c; org[ax] = o;
        ebits |= (unsigned)(e + 127) << (8 * ax);
    }
    if (!ok) { atomicExch(bad, 1); return; }
    unsigned qlo[3] = {0u, 0u, 0u}, qhi[3] = {0u, 0u, 0u};
    int child[4];
    // the node's inner children get CONSECUTIVE slots of the next level (one reservation per node): a ray that goes on into two of
    // them finds both within 256 bytes instead of anywhere in the level
#ifndef W4_CONTIG
#define W4_CONTIG 1
#endif
    int n_inner_ch = 0;
    for (int k = 0; k < cnt; ++k) n_inner_ch += s[k].child >= 0;
    int next_slot = (W4_CONTIG && n_inner_ch) ? atomicAdd(n_out, n_inner_ch) : 0;
    for (int k = 0; k < 4; ++k) {
        if (k >= cnt) {
            child[k] = GPT_W4_EMPTY;
            for (int ax = 0; ax < 3; ++ax) { qlo[ax] |= 255u << (8 * k); }      // inverted box (lo 255, hi 0): entry > exit, never hit
            continue;
        }
        for (int ax = 0; ax < 3; ++ax) {
            const float c = cell[ax], o = org[ax];
            int ql = (int)fminf(fmaxf(floorf((s[k].lo[ax] - o) / c), 1.0f), 254.0f);
            while (ql > 1 && fmaf((float)ql, c, o) > s[k].lo[ax]) --ql;
            int qh = (int)fminf(fmaxf(ceilf((s[k].hi[ax] - o) / c), 1.0f), 254.0f);
            while (qh < 254 && fmaf((float)qh, c, o) < s[k].hi[ax]) ++qh;
            if (fmaf((float)ql, c, o) > s[k].lo[ax] || fmaf((float)qh, c, o) < s[k].hi[ax]) atomicExch(bad, 2);   // cannot happen
            qlo[ax] |= (unsigned)(ql - 1) << (8 * k);
            qhi[ax] |= (unsigned)(qh + 1) << (8 * k);
        }
        if (s[k].child >= 0) {
            const int j = W4_CONTIG ? next_slot++ : atomicAdd(n_out, 1);
            q_out[j] = s[k].child;
            child[k] = (base_next + j) * 4;
        } else {
            child[k] = s[k].child;
        }
    }
    float4* w = wnodes + 4 * (size_t)(base_in + i);
    w[0] = make_float4(org[0], org[1], org[2], __uint_as_float(ebits));
    w[1] = make_float4(__uint_as_float(qlo[0]), __uint_as_float(qhi[0]), __uint_as_float(qlo[1]), __uint_as_float(qhi[1]));
    w[2] = make_float4(__uint_as_float(qlo[2]), __uint_as_float(qhi[2]), 0.0f, 0.0f);
    w[3] = make_float4(__int_as_float(child[0]), __int_as_float(child[1]), __int_as_float(child[2]), __int_as_float(child[3]));
}

cudaError_t gptk_build_wide(const float4* nodes2, int n_inner, WideResult* out, cudaStream_t st)
{
    out->wnodes = nullptr; out->n_wide = 0; out->levels = 0; out->status = 0; out->launches = 0;
    if (n_inner <= 0 || !nodes2) return cudaSuccess;
    float4* tmp = nullptr; int* q[2] = {nullptr, nullptr}; int* ctr = nullptr;
    cudaError_t e;
#define CK(x) do { e = (x); if (e != cudaSuccess) goto fail; } while (0)
    CK(cudaMalloc(&tmp, sizeof(float4) * 4 * (size_t)n_inner));        // at most one wide node per binary node
    CK(cudaMalloc(&q[0], sizeof(int) * (size_t)n_inner));
    CK(cudaMalloc(&q[1], sizeof(int) * (size_t)n_inner));
    CK(cudaMalloc(&ctr, 2 * sizeof(int)));
    CK(cudaMemsetAsync(ctr, 0, 2 * sizeof(int), st));
    CK(cudaMemsetAsync(q[0], 0, sizeof(int), st));                       // level 0 = the binary root (float4 index 0)
    {
        int n_in = 1, base_in = 0, cur = 0, levels = 0;
        while (n_in > 0) {
            const int base_next = base_in + n_in;
            int h[2] = {0, 0};
            k_w4_collapse_level<<<(n_in + 127) / 128, 128, 0, st>>>(nodes2, q[cur], n_in, base_in, base_next, q[cur ^ 1], ctr, tmp, ctr + 1);
            CK(cudaGetLastError());
            ++out->launches;
            CK(cudaMemcpyAsync(h, ctr, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            ++levels;
            if (h[1]) { out->status = h[1]; break; }
            if (base_next + h[0] > n_inner || levels > 256) { out->status = 3; break; }      // not a tree
            CK(cudaMemsetAsync(ctr, 0, sizeof(int), st));
            base_in = base_next; n_in = h[0]; cur ^= 1;
        }
        if (out->status == 0) {
            out->n_wide = base_in; out->levels = levels;
            CK(cudaMalloc(&out->wnodes, sizeof(float4) * 4 * (size_t)out->n_wide));
            CK(cudaMemcpyAsync(out->wnodes, tmp, sizeof(float4) * 4 * (size_t)out->n_wide, cudaMemcpyDeviceToDevice, st));
            CK(cudaStreamSynchronize(st));
        }
    }
    cudaFree(tmp); cudaFree(q[0]); cudaFree(q[1]); cudaFree(ctr);
    return cudaSuccess;
fail:
    cudaFree(tmp); cudaFree(q[0]); cudaFree(q[1]); cudaFree(ctr);
    if (out->wnodes) { cudaFree(out->wnodes); out->wnodes = nullptr; }
    return e;
#undef CK
}
