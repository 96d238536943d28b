// pt_shade.cuh -- camera-ray generation and BSDF/shading arithmetic (sm_100a), in the exact
// operation order of the reference's device code (see pt_math.cuh for the method).
#pragma once
#include "pt_math.cuh"

namespace ptk {

struct PathState {
    V3 o, d, c;          // ray origin, ray direction, throughput colour
    int pixel;
    int remaining;
    int bounces;
};

struct CamRec {           // the fields of struct Camera the kernels use
    int W, H;
    float pos[3], view[3], up[3], right[3];
    float plx, ply;
};

struct OptRec {           // class Settings (pathtrace.cu:31-42)
    int rr, ssaa, dof;
    float focal, aperture;
};

// generateRayFromCamera (pathtrace.cu:184-227); index = x + y*W; iter is 1-based.
__device__ __forceinline__ PathState make_camera_path(const CamRec& cam, const OptRec& opt, int iter,
                                                      int depth, int x, int y)
{
    PathState p;
    const int index = x + y * cam.W;
    p.o = v3(cam.pos[0], cam.pos[1], cam.pos[2]);
    p.c = v3(1.f, 1.f, 1.f);
    uint32_t rng = rng_seed(iter, index, depth);

    float sx = __fsub_rn((float)x, __fmul_rn((float)cam.W, 0.5f));
    float sy = __fsub_rn((float)y, __fmul_rn((float)cam.H, 0.5f));
    if (opt.ssaa) {                       // first draw jitters x, second y (nvcc 12.9 order)
        sx = __fadd_rn(sx, rng_u01(rng));
        sy = __fadd_rn(sy, rng_u01(rng));
    }
    const V3 rp = v3(__fmul_rn(cam.right[0], cam.plx), __fmul_rn(cam.right[1], cam.plx), __fmul_rn(cam.right[2], cam.plx));
    const V3 upv = v3(__fmul_rn(cam.up[0], cam.ply), __fmul_rn(cam.up[1], cam.ply), __fmul_rn(cam.up[2], cam.ply));
    // SASS: both subtractions are fused (FFMA -rp, sx, view ; FFMA -upv, sy, .)
    V3 dir = v3(__fmaf_rn(-upv.x, sy, __fmaf_rn(-rp.x, sx, cam.view[0])),
                __fmaf_rn(-upv.y, sy, __fmaf_rn(-rp.y, sx, cam.view[1])),
                __fmaf_rn(-upv.z, sy, __fmaf_rn(-rp.z, sx, cam.view[2])));
    p.d = normalize_exact(dir);

    if (opt.dof) {                        // thin lens: r = u*aperture (not sqrt), lens in world XY
        const float r = __fmul_rn(rng_u01(rng), opt.aperture);
        const float u = rng_u01(rng);
        const float theta = __fmul_rn(__fadd_rn(u, u), 3.1415926535897932384626422832795028841971f);
        const float ct = cosf(theta), st = sinf(theta);
        const float ft = __fdiv_rn(opt.focal, fabsf(p.d.z));
        const V3 focus = v3(__fmaf_rn(p.d.x, ft, p.o.x), __fmaf_rn(p.d.y, ft, p.o.y), __fmaf_rn(ft, p.d.z, p.o.z));
        p.o = v3(__fmaf_rn(r, ct, p.o.x), __fmaf_rn(r, st, p.o.y), __fadd_rn(p.o.z, 0.0f));
        p.d = normalize_exact(v3(__fsub_rn(focus.x, p.o.x), __fsub_rn(focus.y, p.o.y), __fsub_rn(focus.z, p.o.z)));
    }
    p.pixel = index;
    p.remaining = depth;
    p.bounces = 0;
    return p;
}

// glm::reflect(I, N) = I - N * dot(N, I) * 2, contracted: I - fma(N, d, N*d)
__device__ __forceinline__ V3 reflect_exact(const V3& I, const V3& N)
{
    const float d = dot_exact(N, I);
    return v3(__fsub_rn(I.x, __fmaf_rn(N.x, d, __fmul_rn(N.x, d))),
              __fsub_rn(I.y, __fmaf_rn(N.y, d, __fmul_rn(N.y, d))),
              __fsub_rn(I.z, __fmaf_rn(N.z, d, __fmul_rn(N.z, d))));
}

// calculateRandomDirectionInHemisphere (interactions.cu:3-41)
__device__ __forceinline__ V3 cosine_hemisphere(const V3& n, uint32_t& rng)
{
    const float up = __fsqrt_rn(rng_u01(rng));
    const float around = __fmul_rn(rng_u01(rng), 6.2831853071795864769252867665590057683943f);
    const float over = __fsqrt_rn(__fmaf_rn(-up, up, 1.0f));          // SASS: FFMA -up, up, 1
    const float S = 0.5773502691896257645091487805019574556476f;
    V3 nn;
    if (fabsf(n.x) < S) nn = v3(1.f, 0.f, 0.f);
    else if (fabsf(n.y) < S) nn = v3(0.f, 1.f, 0.f);
    else nn = v3(0.f, 0.f, 1.f);
    const V3 p1 = normalize_exact(cross_exact(n, nn));
    const V3 p2 = normalize_exact(cross_exact(n, p1));
    const float oc = __fmul_rn(over, cosf(around));
    const float os = __fmul_rn(over, sinf(around));
    return v3(__fmaf_rn(p2.x, os, __fmaf_rn(p1.x, oc, __fmul_rn(up, n.x))),
              __fmaf_rn(p2.y, os, __fmaf_rn(p1.y, oc, __fmul_rn(up, n.y))),
              __fmaf_rn(p2.z, os, __fmaf_rn(p1.z, oc, __fmul_rn(up, n.z))));
}

// scatterRay (interactions.cu:43-85).  `base` is the first colour factor (interactions.cu:60): the
// texel of a textured material, m.color otherwise.
__device__ __forceinline__ void scatter_exact(PathState& p, const V3& hit, const V3& n, const MatRec& m,
                                              const V3& base, uint32_t& rng)
{
    const V3 I = p.d;
    p.o = v3(__fmaf_rn(n.x, 1.0e-4f, hit.x), __fmaf_rn(n.y, 1.0e-4f, hit.y), __fmaf_rn(n.z, 1.0e-4f, hit.z));
    p.c = v3(__fmul_rn(base.x, p.c.x), __fmul_rn(base.y, p.c.y), __fmul_rn(base.z, p.c.z));
    if (m.has_refractive != 0.0f) {
        const float ior = m.ior;
        const float a = __fadd_rn(ior, -1.0f), b = __fadd_rn(ior, 1.0f);
        const float R0 = __fdiv_rn(__fmul_rn(a, a), __fmul_rn(b, b));
        const float dn = dot_exact(I, n);
        const float X = __fsub_rn(1.0f, fabsf(dn));
        const float X2 = __fmul_rn(X, X);
        const float R = __fmaf_rn(__fsub_rn(1.0f, R0), __fmul_rn(X2, __fmul_rn(X, X2)), R0);
        const float u = rng_u01(rng);
        if (R < u) {                      // glm::refract(I, n, ior): eta is the ior itself
            // SASS: k = FFMA(-(ior*ior), FFMA(-dn, dn, 1), 1); d = FFMA(n, -s, ior*I) * sel
            const float k = __fmaf_rn(-__fmul_rn(ior, ior), __fmaf_rn(-dn, dn, 1.0f), 1.0f);
            const float s = __fmaf_rn(ior, dn, __fsqrt_rn(k));
            const float sel = k >= 0.0f ? 1.0f : 0.0f;
            p.d = v3(__fmul_rn(__fmaf_rn(n.x, -s, __fmul_rn(ior, I.x)), sel),
                     __fmul_rn(__fmaf_rn(n.y, -s, __fmul_rn(ior, I.y)), sel),
                     __fmul_rn(__fmaf_rn(n.z, -s, __fmul_rn(ior, I.z)), sel));
            p.c = v3(__fmul_rn(m.color[0], p.c.x), __fmul_rn(m.color[1], p.c.y), __fmul_rn(m.color[2], p.c.z));
        } else {
            p.d = reflect_exact(I, n);
            p.c = v3(__fmul_rn(m.spec_color[0], p.c.x), __fmul_rn(m.spec_color[1], p.c.y), __fmul_rn(m.spec_color[2], p.c.z));
        }
    } else if (rng_u01(rng) < m.has_reflective) {
        p.d = reflect_exact(I, n);
        p.c = v3(__fmul_rn(m.spec_color[0], p.c.x), __fmul_rn(m.spec_color[1], p.c.y), __fmul_rn(m.spec_color[2], p.c.z));
    } else {
        p.d = cosine_hemisphere(n, rng);
    }
    p.c = v3(__fmul_rn(m.color[0], p.c.x), __fmul_rn(m.color[1], p.c.y), __fmul_rn(m.color[2], p.c.z));
}

__device__ __forceinline__ bool shade_exact(PathState& p, float t, const V3& n, const MatRec& m, int iter,
                                            int idx, int rr);

// shadeMaterials (pathtrace.cu:301-344) for one path at compacted slot `idx`.
// Returns true when the path stays alive; on return p.c is the colour finalGather would add
// for a terminated path (0 unless it ended on a light).
__device__ __forceinline__ bool shade_exact(PathState& p, float t, const V3& n, const MatRec& m, const V3& base,
                                            int iter, int idx, int rr)
{
    if (!(t > 0.0f)) {
        p.c = v3(0.f, 0.f, 0.f);
        p.remaining = 0;
        return false;
    }
    uint32_t rng = rng_seed(iter, idx, p.remaining);
    if (m.emittance > 0.0f) {
        p.c = v3(__fmul_rn(__fmul_rn(m.color[0], m.emittance), p.c.x),
                 __fmul_rn(__fmul_rn(m.color[1], m.emittance), p.c.y),
                 __fmul_rn(__fmul_rn(m.color[2], m.emittance), p.c.z));
        p.remaining = 0;
        return false;
    }
    const V3 hit = point_on_ray(p.o, p.d, t);
    scatter_exact(p, hit, n, m, base, rng);
    const int old_bounces = p.bounces;
    p.bounces = old_bounces + 1;
    p.remaining -= 1;
    if (p.remaining == 0) {
        p.c = v3(0.f, 0.f, 0.f);
        return false;
    }
    if (rr && old_bounces >= 3) {
        const float luma = __fmaf_rn(p.c.z, 0.0722f, __fmaf_rn(p.c.x, 0.2126f, __fmul_rn(p.c.y, 0.7152f)));
        const float x = __fsub_rn(1.0f, luma);
        const float q = x < 0.05f ? 0.05f : x;
        const float u = rng_u01(rng);
        if (u < q) {
            p.c = v3(0.f, 0.f, 0.f);
            p.remaining = 0;
            return false;
        }
        const float inv = __fsub_rn(1.0f, q);
        p.c = v3(__fdiv_rn(p.c.x, inv), __fdiv_rn(p.c.y, inv), __fdiv_rn(p.c.z, inv));
    }
    return true;
}

// texture-less convenience form
__device__ __forceinline__ bool shade_exact(PathState& p, float t, const V3& n, const MatRec& m, int iter,
                                            int idx, int rr)
{
    return shade_exact(p, t, n, m, v3(m.color[0], m.color[1], m.color[2]), iter, idx, rr);
}

}  // namespace ptk
