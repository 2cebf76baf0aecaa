/*
 * oracle/oracle_math.h -- TEST INFRASTRUCTURE ONLY (see oracle.h).
 *
 * Scalar vector helpers, the counter-based RNG and the deterministic transcendental
 * functions of the "numeric contract" (DESIGN.md).  Replaces, on the oracle side:
 *   glm::dot/cross/normalize as used in tracer.cu / cudaUtils.h,
 *   curand_init/curand_uniform (tracer.cu:362-363; cudaUtils.h:117-118),
 *   cosf/sinf (cudaUtils.h:190, tracer.cu:280) and powf (tracer.cu:267).
 * Philox4x32-10 is Salmon et al., "Parallel random numbers: as easy as 1, 2, 3" (SC'11);
 * it is pinned by the published Random123 known-answer vectors in tests/test_oracle_kat.py.
 */
#ifndef ORACLE_MATH_H
#define ORACLE_MATH_H
#include <math.h>
#include <stdint.h>
#include <string.h>

struct V3 { float x, y, z; };
static inline V3 v3(float x, float y, float z) { V3 r = {x, y, z}; return r; }
static inline V3 vadd(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline V3 vsub(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline V3 vmul(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline V3 vscale(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
/* a + b*s with one rounding per component */
static inline V3 vmad(V3 b, float s, V3 a) { return v3(fmaf(b.x, s, a.x), fmaf(b.y, s, a.y), fmaf(b.z, s, a.z)); }
/* glm::dot = (a.x*b.x + a.y*b.y) + a.z*b.z as nvcc contracts it in the reference kernel (read off its SASS, e.g. the
 * determinant of cudaUtils.h:145: FMUL y, FFMA x, FFMA z): the y product is rounded, x and z are fused */
static inline float vdot(V3 a, V3 b) { return fmaf(a.z, b.z, fmaf(a.x, b.x, a.y * b.y)); }
static inline V3 vcross(V3 a, V3 b)
{
    return v3(fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x)));
}
/* glm::normalize = v * inversesqrt(dot(v,v)), inversesqrt = 1/sqrt (IEEE div + sqrt) */
static inline V3 vnormalize(V3 a) { float inv = 1.0f / sqrtf(vdot(a, a)); return vscale(a, inv); }

static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

/* ---- Philox4x32-10 ---- */
static inline void philox4x32_10(const uint32_t c_in[4], const uint32_t k_in[2], uint32_t out[4])
{
    uint32_t c0 = c_in[0], c1 = c_in[1], c2 = c_in[2], c3 = c_in[3];
    uint32_t k0 = k_in[0], k1 = k_in[1];
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* 24 high bits -> {2^-24 .. 1}: the (0,1] range contract of curand_uniform */
static inline float u01(uint32_t x) { return (float)((x >> 8) + 1u) * 5.9604644775390625e-8f; }

/* RNG block for (pixel, global sample index, dimension block): block 0 = camera jitter,
 * block d+1 = bounce d.  key = 64-bit seed. */
static inline void rng_block(uint64_t seed, uint32_t pixel, uint64_t sample, uint32_t block, float u[4])
{
    uint32_t c[4] = { pixel, (uint32_t)sample, (uint32_t)(sample >> 32), block };
    uint32_t k[2] = { (uint32_t)seed, (uint32_t)(seed >> 32) };
    uint32_t o[4];
    philox4x32_10(c, k, o);
    u[0] = u01(o[0]); u[1] = u01(o[1]); u[2] = u01(o[2]); u[3] = u01(o[3]);
}

/* sin/cos of 2*pi*u, u in [0,1]: quadrant reduction (exact) + Taylor polynomials on
 * |x| <= pi/4 evaluated with FMAs only, so CPU and GPU agree bit for bit. */
static inline void sincos2pi(float u, float* s_out, float* c_out)
{
    float q = rintf(u * 4.0f);
    float r = fmaf(q, -0.25f, u);              /* exact, |r| <= 1/8 */
    float x = r * 6.2831853071795864769f;
    float x2 = x * x;
    float sp = fmaf(x2, 2.7557319223985893e-6f, -1.9841269841269841e-4f);
    sp = fmaf(sp, x2, 8.3333333333333333e-3f);
    sp = fmaf(sp, x2, -1.6666666666666667e-1f);
    float s = fmaf(x * x2, sp, x);
    float cp = fmaf(x2, -2.7557319223985888e-7f, 2.4801587301587302e-5f);
    cp = fmaf(cp, x2, -1.3888888888888889e-3f);
    cp = fmaf(cp, x2, 4.1666666666666664e-2f);
    cp = fmaf(cp, x2, -0.5f);
    float c = fmaf(x2, cp, 1.0f);
    switch (((int)q) & 3) {
        case 0: *s_out = s;  *c_out = c;  break;
        case 1: *s_out = c;  *c_out = -s; break;
        case 2: *s_out = -s; *c_out = -c; break;
        default: *s_out = -c; *c_out = s; break;
    }
}

/* x^y for x >= 0 through log2/exp2 polynomials (FMA-only, bit-reproducible).
 * Used for the Phong lobe cosTheta = (1-r2)^(1/(phongExpo+1)) (tracer.cu:267). */
static inline float det_powf(float x, float y)
{
    if (!(x > 1.17549435e-38f)) return 0.0f;           /* 0, denormal, negative, NaN -> 0 */
    uint32_t b = f2u(x);
    int e = (int)((b >> 23) & 255u) - 127;
    float m = u2f((b & 0x007fffffu) | 0x3f800000u);    /* [1,2) */
    if (m > 1.41421356f) { m *= 0.5f; e += 1; }
    float s = (m - 1.0f) / (m + 1.0f);
    float s2 = s * s;
    float p = fmaf(s2, 9.0909090909090912e-2f, 1.1111111111111111e-1f);
    p = fmaf(p, s2, 1.4285714285714285e-1f);
    p = fmaf(p, s2, 0.2f);
    p = fmaf(p, s2, 3.3333333333333331e-1f);
    p = fmaf(p, s2, 1.0f);
    float lnm = 2.0f * s * p;
    float l2 = fmaf(lnm, 1.4426950408889634f, (float)e);
    float z = y * l2;
    if (!(z > -125.0f)) return 0.0f;
    if (z > 127.0f) z = 127.0f;
    float zi = rintf(z);
    float g = (z - zi) * 0.69314718055994531f;
    float q = fmaf(g, 1.9841269841269841e-4f, 1.3888888888888889e-3f);
    q = fmaf(q, g, 8.3333333333333333e-3f);
    q = fmaf(q, g, 4.1666666666666664e-2f);
    q = fmaf(q, g, 1.6666666666666667e-1f);
    q = fmaf(q, g, 0.5f);
    q = fmaf(q, g, 1.0f);
    q = fmaf(q, g, 1.0f);
    return u2f(f2u(q) + ((uint32_t)(int)zi << 23));
}

#endif
