// gpt_math.cuh -- device math of the numeric contract (DESIGN.md "Numeric contract").
//
// The translation unit is compiled with -fmad=false: the ONLY fused multiply-adds are the
// explicit fmaf() calls below and in the traversal/shading code, so results are bit-reproducible
// against a scalar CPU evaluation of the same formulas.  Division and sqrt are the IEEE-rounded
// forms (nvcc defaults -prec-div=true -prec-sqrt=true).
//
// Replaces, on the device side: glm vec3 ops compiled as device code (tracer.cu, cudaUtils.h),
// cuRAND XORWOW (tracer.cu:362-363) -> Philox4x32-10 counter-based RNG (north_star item 3),
// cosf/sinf/powf (cudaUtils.h:190, tracer.cu:267,280) -> deterministic polynomial forms.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define GPT_F32_MAX 3.402823466e+38f
#define GPT_FULL_MASK 0xffffffffu

struct f3 { float x, y, z; };

// One 64-byte record (a BVH node) as TWO 256-bit loads through the read-only path (LDG.E.256, new on sm_100) instead of four
// 128-bit ones: a divergent warp costs the L1 one tag look-up ("wavefront") per lane and per load INSTRUCTION, and the extend
// kernel runs at ~0.9 of the rate at which L1 serves such gathers (bench.py roofline.bounds), so halving the loads per node
// halves that cost.  GPT_LD256=0 keeps the four LDG.128 (A/B).  p must be 32-byte aligned (nodes are 64-byte aligned).
#ifndef GPT_LD256
#define GPT_LD256 1
#endif
__device__ __forceinline__ void ldg_node64(const float4* __restrict__ p, float4& a, float4& b, float4& c, float4& d)
{
#if GPT_LD256
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p));
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(c.x), "=f"(c.y), "=f"(c.z), "=f"(c.w), "=f"(d.x), "=f"(d.y), "=f"(d.z), "=f"(d.w) : "l"(p + 2));
#else
    a = __ldg(p); b = __ldg(p + 1); c = __ldg(p + 2); d = __ldg(p + 3);
#endif
}

__device__ __forceinline__ f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ f3 add3(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ f3 sub3(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ f3 mul3(f3 a, f3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ f3 scale3(f3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
// a + b*s, one rounding per component
__device__ __forceinline__ f3 mad3(f3 b, float s, f3 a) { return mk3(fmaf(b.x, s, a.x), fmaf(b.y, s, a.y), fmaf(b.z, s, a.z)); }
// glm::dot as nvcc contracts it in the reference's kernel (read off its SASS): the y product rounded, x and z fused onto it
__device__ __forceinline__ float dot3(f3 a, f3 b) { return fmaf(a.z, b.z, fmaf(a.x, b.x, a.y * b.y)); }
__device__ __forceinline__ f3 cross3(f3 a, f3 b)
{
    return mk3(fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x)));
}
__device__ __forceinline__ f3 normalize3(f3 a) { float inv = 1.0f / sqrtf(dot3(a, a)); return scale3(a, inv); }

// ---- Philox4x32-10 (Salmon et al., SC'11) ----
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k)
{
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u; k.y += 0xBB67AE85u;
    }
    return c;
}

// 24 high bits -> {2^-24 .. 1}: keeps curand_uniform's (0,1] range contract
__device__ __forceinline__ float u01(uint32_t x) { return (float)((x >> 8) + 1u) * 5.9604644775390625e-8f; }

// RNG block keyed by (seed; pixel, global sample index, dimension block).  Block 0 = camera jitter,
// block d+1 = bounce d.
__device__ __forceinline__ float4 rng_block(uint64_t seed, uint32_t pixel, uint64_t sample, uint32_t block)
{
    const uint4 o = philox4x32_10(make_uint4(pixel, (uint32_t)sample, (uint32_t)(sample >> 32), block),
                                  make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
    return make_float4(u01(o.x), u01(o.y), u01(o.z), u01(o.w));
}

// sin/cos(2*pi*u), u in [0,1]: exact quadrant reduction + FMA-only Taylor polynomials on |x|<=pi/4
__device__ __forceinline__ void sincos2pi(float u, float& s_out, float& c_out)
{
    const float q = rintf(u * 4.0f);
    const float r = fmaf(q, -0.25f, u);
    const float x = r * 6.2831853071795864769f;
    const float x2 = x * x;
    float sp = fmaf(x2, 2.7557319223985893e-6f, -1.9841269841269841e-4f);
    sp = fmaf(sp, x2, 8.3333333333333333e-3f);
    sp = fmaf(sp, x2, -1.6666666666666667e-1f);
    const float s = fmaf(x * x2, sp, x);
    float cp = fmaf(x2, -2.7557319223985888e-7f, 2.4801587301587302e-5f);
    cp = fmaf(cp, x2, -1.3888888888888889e-3f);
    cp = fmaf(cp, x2, 4.1666666666666664e-2f);
    cp = fmaf(cp, x2, -0.5f);
    const float c = fmaf(x2, cp, 1.0f);
    const int qi = ((int)q) & 3;
    const float a = (qi & 1) ? c : s;       // |sin| source
    const float b = (qi & 1) ? s : c;       // |cos| source
    s_out = (qi & 2) ? -a : a;
    c_out = (qi == 1 || qi == 2) ? -b : b;
}

// x^y, x >= 0, via FMA-only log2/exp2 polynomials (Phong lobe, tracer.cu:267)
__device__ __forceinline__ float det_powf(float x, float y)
{
    if (!(x > 1.17549435e-38f)) return 0.0f;
    const uint32_t b = __float_as_uint(x);
    int e = (int)((b >> 23) & 255u) - 127;
    float m = __uint_as_float((b & 0x007fffffu) | 0x3f800000u);
    if (m > 1.41421356f) { m *= 0.5f; e += 1; }
    const float s = (m - 1.0f) / (m + 1.0f);
    const float s2 = s * s;
    float p = fmaf(s2, 9.0909090909090912e-2f, 1.1111111111111111e-1f);
    p = fmaf(p, s2, 1.4285714285714285e-1f);
    p = fmaf(p, s2, 0.2f);
    p = fmaf(p, s2, 3.3333333333333331e-1f);
    p = fmaf(p, s2, 1.0f);
    const float lnm = 2.0f * s * p;
    const float l2 = fmaf(lnm, 1.4426950408889634f, (float)e);
    float z = y * l2;
    if (!(z > -125.0f)) return 0.0f;
    if (z > 127.0f) z = 127.0f;
    const float zi = rintf(z);
    const float g = (z - zi) * 0.69314718055994531f;
    float q = fmaf(g, 1.9841269841269841e-4f, 1.3888888888888889e-3f);
    q = fmaf(q, g, 8.3333333333333333e-3f);
    q = fmaf(q, g, 4.1666666666666664e-2f);
    q = fmaf(q, g, 1.6666666666666667e-1f);
    q = fmaf(q, g, 0.5f);
    q = fmaf(q, g, 1.0f);
    q = fmaf(q, g, 1.0f);
    return __uint_as_float(__float_as_uint(q) + ((uint32_t)(int)zi << 23));
}


// ---- GPT_FLAG_FAST arithmetic (opt-in; held to north_star's statistical tolerance, NOT bit-exact to the oracle) ------------
// The compat contract pins IEEE sqrt + divide normalisations and FMA-only sin/cos/pow polynomials so that the CPU oracle can be
// matched bit for bit.  FAST swaps them for the special-function unit: MUFU.RSQ / RCP / SIN / COS / EX2 / LG2 (1-2 ulp), and lets
// the remaining products contract.  The traversal (nearest hit, face index) is NOT affected: only shading arithmetic changes.
template <bool FAST> __device__ __forceinline__ f3 norm3_t(f3 a)
{
    if (FAST) { const float inv = rsqrtf(dot3(a, a)); return scale3(a, inv); }
    return normalize3(a);
}
template <bool FAST> __device__ __forceinline__ float sqrt_t(float x) { return FAST ? __fmul_rn(x, rsqrtf(fmaxf(x, 1e-30f))) : sqrtf(x); }
template <bool FAST> __device__ __forceinline__ float fdiv_t(float a, float b) { return FAST ? __fdividef(a, b) : a / b; }
template <bool FAST> __device__ __forceinline__ void sincos2pi_t(float u, float& s, float& c)
{
    if (FAST) { __sincosf(6.2831853071795864769f * (u - rintf(u)), &s, &c); }      // argument reduced to [-pi, pi] for MUFU.SIN / COS
    else sincos2pi(u, s, c);
}
template <bool FAST> __device__ __forceinline__ float pow_t(float x, float y) { return FAST ? __powf(x, y) : det_powf(x, y); }
