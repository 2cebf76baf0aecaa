// gpt_shade.cuh -- camera ray generation and the per-bounce shading step.
//
// Replaces getCamRayDir (cudaUtils.h:111-134), uniformSampleHemisphere (:185-192) and the body of the
// bounce loop of getSample (tracer.cu:72-300), including every quirk SURVEY.md 8-Q lists (Q1 METAL `w`
// = image width, Q2 unweighted miss colour, Q4 un-flipped triangle normal, Q5 non-cosine diffuse lobe,
// Q7 REFR constants).  One call = one bounce, so the same code serves the megakernel and the wavefront
// shade kernel.
#pragma once
#include "gpt_traverse.cuh"

struct PathState {
    f3 o, d;        // current ray
    f3 mask;        // colour mask   (tracer.cu:68)
    f3 accu;        // accumulated colour (tracer.cu:69)
};

// getCamRayDir, cudaUtils.h:111-134.  Row 0 is the bottom of the image; origin lies ON the image plane.
__device__ __forceinline__ void camera_ray(const RenderParams& p, int px, int py, float jx, float jy, f3& o, f3& d)
{
    float xs = ((float)px - (float)p.width / 2.0f) + 0.5f;
    xs = xs + jx;
    xs = xs * p.dist; xs = xs * p.aspect; xs = xs * p.fov; xs = xs / (float)(p.width - 1);
    float ys = ((float)py - (float)p.height / 2.0f) + 0.5f;
    ys = ys + jy;
    ys = ys * p.dist; ys = ys * p.fov; ys = ys / (float)(p.height - 1);
    const f3 dir = mad3(ld3(p.up), ys, mad3(ld3(p.front), p.dist, scale3(ld3(p.right), xs)));   // FMUL right, FFMA front, FFMA up (reference SASS)
    o = add3(ld3(p.pos), dir);
    d = normalize3(dir);
}

__device__ __forceinline__ void start_path(const RenderParams& p, int px, int py, uint32_t pixel, uint64_t sample, PathState& ps)
{
    float jx = 0.0f, jy = 0.0f;
    if (!(p.flags & RP_NO_JITTER)) {
        const float4 u = rng_block(p.seed, pixel, sample, 0u);
        jx = u.x - 0.5f; jy = u.y - 0.5f;          // cudaUtils.h:117-118
    }
    camera_ray(p, px, py, jx, jy, ps.o, ps.d);
    ps.mask = mk3(1.0f, 1.0f, 1.0f);
    ps.accu = mk3(0.0f, 0.0f, 0.0f);
}

// Box::normalAt spec, cudaUtils.h:61-76
__device__ __forceinline__ f3 box_normal(const float* bx, f3 p)
{
    const float eps = 0.001f;
    if (fabsf(bx[0] - p.x) < eps) return mk3(-1, 0, 0);
    if (fabsf(bx[3] - p.x) < eps) return mk3(1, 0, 0);
    if (fabsf(bx[1] - p.y) < eps) return mk3(0, -1, 0);
    if (fabsf(bx[4] - p.y) < eps) return mk3(0, 1, 0);
    if (fabsf(bx[2] - p.z) < eps) return mk3(0, 0, -1);
    return mk3(0, 0, 1);
}

// One bounce of tracer.cu:112-296.  Returns false when the path ended on a miss (ps.accu then holds
// the background colour, Q2); true when the path continues with the new ray in ps.
// accu_slot != nullptr (wavefront pipeline): the accumulated colour lives in memory and ps.accu is not used.  It changes
// only on a miss (overwritten with the background, no read) and on an emitting hit (read-modify-write); on any other
// hit fma(mask, 0, accu) == accu bit for bit for a finite mask, so the slot is neither read nor written.
// FIRST (bounce 0 of the wavefront pipeline): the accumulated colour is still 0 (tracer.cu:69) and the slot holds nothing yet:
// it is written unconditionally, never read.
// FAST: GPT_FLAG_FAST arithmetic (gpt_math.cuh): same formulas, special-function unit instead of the IEEE / polynomial forms.
template <bool FIRST = false, bool FAST = false>
__device__ __forceinline__ bool shade_bounce(const DevScene& sc, const RenderParams& p,
                                             const float* s_spheres, const float* s_boxes,
                                             const SceneHit h, const uint32_t pixel, const uint64_t sample,
                                             const uint32_t bounce, PathState& ps, float4* accu_slot = nullptr)
{
    // RNG first, while the warp is still converged: placed after the geometry branches the compiler
    // tail-duplicates the ten Philox rounds into each branch and the warp runs them once per geometry type
    // (ncu: 18 of 32 lanes active on those instructions).
    const float4 u = rng_block(p.seed, pixel, sample, bounce + 1u);
    if (h.geom == GEO_NONE) {                                            // tracer.cu:140-142
        if (accu_slot) *accu_slot = make_float4(p.bk[0], p.bk[1], p.bk[2], 0.0f); else ps.accu = ld3(p.bk);
        return false;
    }
    f3 hitpos = mad3(ps.d, h.t, ps.o);                                   // :112
    // The geometry branches only gather the un-normalised normal; the normalisation (an IEEE square root and a division) and the
    // face-forwarding run once, after the branches have re-converged, instead of once per geometry type present in the warp.
    f3 nr, objcol, emit; int mat;
    bool unit = false, face_forward = true;
    if (h.geom == GEO_SPHERE) {                                          // :113-122
        const float* s = s_spheres + 11 * h.prim;
        nr = sub3(hitpos, mk3(s[0], s[1], s[2]));
        emit = mk3(s[4], s[5], s[6]); objcol = mk3(s[7], s[8], s[9]); mat = __float_as_int(s[10]);
    } else if (h.geom == GEO_TRI) {                                      // :123-139
        float4 b, c;
        tri_rows_bc(sc.tris, h.prim, b, c);
        // cross(v0-v1, v0-v2) == cross(e1, e2) bit for bit (negation is exact)   cudaUtils.h:432
        nr = cross3(mk3(b.x, b.y, b.z), mk3(c.x, c.y, c.z));
        face_forward = false;                                            // Q4: not face-forwarded
        emit = ld3(p.emi); objcol = ld3(p.col); mat = p.tri_mat;
        if (sc.n_faces) {                                                // N4: per-face material table
            const int face = __float_as_int(tri_row_a(sc.tris, h.prim).w);
            if ((unsigned)face < (unsigned)sc.n_faces) {
                const int m = __ldg(sc.face_mat + face);
                if ((unsigned)m < (unsigned)sc.n_tri_mats) {
                    const float* tm = sc.tri_mats + 7 * (size_t)m;
                    objcol = mk3(__ldg(tm), __ldg(tm + 1), __ldg(tm + 2)); emit = mk3(__ldg(tm + 3), __ldg(tm + 4), __ldg(tm + 5));
                    mat = __float_as_int(__ldg(tm + 6));
                }
            }
        }
    } else {                                                             // BOX (commented branch :143-152)
        const float* bx = s_boxes + 13 * h.prim;
        nr = box_normal(bx, hitpos); unit = true;                        // an axis: used as it is
        emit = mk3(bx[6], bx[7], bx[8]); objcol = mk3(bx[9], bx[10], bx[11]); mat = __float_as_int(bx[12]);
    }
    f3 n = norm3_t<FAST>(nr);
    if (unit) n = nr;
    const f3 nl = (face_forward && !(dot3(n, ps.d) < 0.0f)) ? scale3(n, -1.0f) : n;
    if (FIRST && accu_slot) {
        *accu_slot = make_float4(fmaf(ps.mask.x, emit.x, 0.0f), fmaf(ps.mask.y, emit.y, 0.0f), fmaf(ps.mask.z, emit.z, 0.0f), 0.0f);
    } else if (accu_slot) {
        const bool touch = emit.x != 0.0f || emit.y != 0.0f || emit.z != 0.0f
                           || !(fabsf(ps.mask.x) <= GPT_F32_MAX) || !(fabsf(ps.mask.y) <= GPT_F32_MAX) || !(fabsf(ps.mask.z) <= GPT_F32_MAX);
        if (touch) {
            const float4 a = *accu_slot;
            *accu_slot = make_float4(fmaf(ps.mask.x, emit.x, a.x), fmaf(ps.mask.y, emit.y, a.y), fmaf(ps.mask.z, emit.z, a.z), 0.0f);
        }
    } else
    ps.accu = mk3(fmaf(ps.mask.x, emit.x, ps.accu.x), fmaf(ps.mask.y, emit.y, ps.accu.y), fmaf(ps.mask.z, emit.z, ps.accu.z));   // :120/:133, contracted

    f3 nextdir, mscale = objcol;
    float eps = 0.001f;
    if (mat == MAT_DIFF) {                                               // :156-186 (Q5)
        f3 nt = (fabsf(nl.x) > fabsf(nl.y)) ? mk3(nl.z, 0.0f, -nl.x) : mk3(0.0f, -nl.z, nl.y);
        nt = norm3_t<FAST>(nt);
        const f3 nb = norm3_t<FAST>(cross3(nl, nt));
        float sn, cs; sincos2pi_t<FAST>(u.x, sn, cs);
        const float f2 = u.y;                                            // sin(theta), uniform (cudaUtils.h:185-192)
        const f3 rv = mk3(cs * f2, sqrt_t<FAST>(fmaf(-f2, f2, 1.0f)), sn * f2);
        nextdir = mk3(fmaf(nt.x, rv.z, fmaf(nl.x, rv.y, nb.x * rv.x)),
                      fmaf(nt.y, rv.z, fmaf(nl.y, rv.y, nb.y * rv.x)),
                      fmaf(nt.z, rv.z, fmaf(nl.z, rv.y, nb.z * rv.x)));
    } else if (mat == MAT_SPEC) {                                        // :190-203 (Q6)
        // :194 'dir - nl * dot * 2.0f' the way nvcc emits it for the reference: pr = rn(nl*k), 2*pr formed as fma(k, nl, pr), plain subtract
        const float k = dot3(nl, ps.d);
        const f3 pr = scale3(nl, k);
        nextdir = mk3(ps.d.x - fmaf(k, nl.x, pr.x), ps.d.y - fmaf(k, nl.y, pr.y), ps.d.z - fmaf(k, nl.z, pr.z));
    } else if (mat == MAT_REFR) {                                        // :205-256 (Q7)
        const bool into = dot3(n, nl) > 0.0f;
        const float nc = p.air, ntr = p.glass;
        const float nnt = into ? fdiv_t<FAST>(nc, ntr) : fdiv_t<FAST>(ntr, nc);
        const float ddn = dot3(ps.d, nl);
        const float cos2t = fmaf(-(nnt * nnt), fmaf(-ddn, ddn, 1.0f), 1.0f);        // :212, contracted as in the reference's SASS
        mscale = mk3(1.0f, 1.0f, 1.0f);                                  // total internal reflection leaves the mask alone (x * 1 is exact)
        if (cos2t < 0.0f) {
            const float k = dot3(n, ps.d) * 2.0f;
            nextdir = mad3(n, -k, ps.d);
        } else {
            const float s = (into ? 1.0f : -1.0f) * fmaf(ddn, nnt, sqrt_t<FAST>(cos2t));
            const f3 tdir = norm3_t<FAST>(mad3(n, -s, scale3(ps.d, nnt)));
            const float R0 = ((ntr - nc) * (ntr - nc)) / (ntr + nc) * (ntr + nc);   // :230 precedence kept
            const float c = 1.0f - (into ? -ddn : dot3(tdir, n));
            const float Re = fmaf((1.0f - R0) * c * c * c * c, c, R0);           // :232, last product fused into the sum
            const float Tr = 1.0f - Re;
            const float P = 0.25f + 0.5f * Re;
            const float RP = fdiv_t<FAST>(Re, P), TP = fdiv_t<FAST>(Tr, 1.0f - P);
            if (u.x < 0.2f) {                                            // :239 constant, not P
                mscale = mk3(RP, RP, RP);
                const float k = dot3(n, ps.d) * 2.0f;
                nextdir = mad3(n, -k, ps.d);
            } else {
                mscale = mk3(TP, TP, TP);
                nextdir = tdir;                                          // normalised a second time below, as the reference does
            }
        }
    } else {                                                             // METAL :257-293 (Q1)
        float sphi, cphi; sincos2pi_t<FAST>(u.x, sphi, cphi);
        const float cosT = pow_t<FAST>(1.0f - u.y, fdiv_t<FAST>(1.0f, p.phong + 1.0f));
        const float sinT = sqrt_t<FAST>(fmaf(-cosT, cosT, 1.0f));
        const float k = dot3(nl, ps.d) * 2.0f;
        const f3 w1 = norm3_t<FAST>(mad3(nl, -k, ps.d));
        const f3 uu = norm3_t<FAST>(cross3(fabsf(w1.x) >= 0.1f ? mk3(0, 1, 0) : mk3(1, 0, 0), w1));
        const f3 vv = cross3(w1, uu);
        const float a = cphi * sinT, b = sphi * sinT;
        if (p.flags & RP_METAL_FIXED) {
            nextdir = mk3(fmaf(w1.x, cosT, fmaf(vv.x, b, uu.x * a)),
                          fmaf(w1.y, cosT, fmaf(vv.y, b, uu.y * a)),
                          fmaf(w1.z, cosT, fmaf(vv.z, b, uu.z * a)));
        } else {
            const float wc = (float)p.width * cosT;                      // :280 `w` = int image width
            nextdir = mk3(fmaf(vv.x, b, uu.x * a) + wc, fmaf(vv.y, b, uu.y * a) + wc, fmaf(vv.z, b, uu.z * a) + wc);
        }
        eps = 0.0001f;
    }
    // common tail of the four materials, after re-convergence: normalise the new direction, lift the origin off the surface, tint
    nextdir = norm3_t<FAST>(nextdir);
    hitpos = mad3(nl, eps, hitpos);
    ps.mask = mul3(ps.mask, mscale);
    ps.o = hitpos; ps.d = nextdir;
    return true;
}
