// pt_math.cuh -- device arithmetic of the wavefront path tracer (sm_100a).
//
// Every floating-point expression here is written with explicit round-to-nearest intrinsics
// (__fmaf_rn / __fmul_rn / __fadd_rn / __fsub_rn / __fdiv_rn / __fsqrt_rn / __frcp_rn) in the
// operation order the reference's device code has after nvcc 12.9's FMA contraction, so that a
// path traced here takes bit-for-bit the same values as in the reference's CUDA build and the
// per-(iteration, slot, depth) seeded random paths are reproduced exactly.
// The intrinsics are never re-associated or contracted, so the result does not depend on this
// file's compiler heuristics.  Operation order per function was taken from the SASS of the
// reference build (nvcc 12.9, sm_100a) for the sources cited at each function -- SASS, not PTX:
// ptxas fuses PTX's unsuffixed mul.f32 + add/sub.f32 pairs into FFMA on its own (e.g. 1 - up*up,
// glm::cross, view - right*sx), and which product of a*b - c*d it fuses varies per site.
//
// vec3 helpers are deliberately tiny: GLM is not used.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace ptk {

struct V3 {
    float x, y, z;
};

__device__ __forceinline__ V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }

// glm::normalize (GLM 0.9.6 func_geometric.inl:153-159: x * inversesqrt(dot(x, x)));
// dot = fma(z, z, fma(x, x, y*y)); inversesqrt = rcp(sqrt(.)).
__device__ __forceinline__ V3 normalize_exact(const V3& v)
{
    const float d = __fmaf_rn(v.z, v.z, __fmaf_rn(v.x, v.x, __fmul_rn(v.y, v.y)));
    const float s = __frcp_rn(__fsqrt_rn(d));
    return v3(__fmul_rn(v.x, s), __fmul_rn(v.y, s), __fmul_rn(v.z, s));
}

// glm::dot(a, b) as contracted by nvcc: fma(a.z, b.z, fma(a.x, b.x, a.y*b.y))
__device__ __forceinline__ float dot_exact(const V3& a, const V3& b)
{
    return __fmaf_rn(a.z, b.z, __fmaf_rn(a.x, b.x, __fmul_rn(a.y, b.y)));
}

// glm::length(a - b): sqrt(dot(d, d))
__device__ __forceinline__ float distance_exact(const V3& a, const V3& b)
{
    const V3 d = v3(__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y), __fsub_rn(a.z, b.z));
    return __fsqrt_rn(__fmaf_rn(d.z, d.z, __fmaf_rn(d.x, d.x, __fmul_rn(d.y, d.y))));
}

// glm::cross as it ends up in the reference's SASS: the second product is rounded, the first
// is fused into the subtraction (FMUL; FFMA a, b, -t).
__device__ __forceinline__ V3 cross_exact(const V3& a, const V3& b)
{
    return v3(__fmaf_rn(a.y, b.z, -__fmul_rn(b.y, a.z)),
              __fmaf_rn(a.z, b.x, -__fmul_rn(b.z, a.x)),
              __fmaf_rn(a.x, b.y, -__fmul_rn(b.x, a.y)));
}

// ---- RNG -----------------------------------------------------------------------------------
// utilhash: path_tracer/src/intersections.h:13-22
__device__ __forceinline__ uint32_t utilhash(uint32_t a)
{
    a = (a + 0x7ed55d16u) + (a << 12);
    a = (a ^ 0xc761c23cu) ^ (a >> 19);
    a = (a + 0x165667b1u) + (a << 5);
    a = (a + 0xd3a2646cu) ^ (a << 9);
    a = (a + 0xfd7046c5u) + (a << 3);
    a = (a ^ 0xb55a4f09u) ^ (a >> 16);
    return a;
}

// makeSeededRandomEngine (pathtrace.cu:57-62) + thrust::minstd_rand seeding
// (state = seed mod (2^31-1), 0 -> 1).
__device__ __forceinline__ uint32_t rng_seed(int iter, int index, int depth)
{
    const uint32_t h = utilhash((uint32_t)((1 << 31) | (depth << 22) | iter)) ^ utilhash((uint32_t)index);
    const uint32_t s = h % 2147483647u;
    return s == 0u ? 1u : s;
}

// one minstd step x <- 48271 x mod (2^31-1) (Schrage, q = 44488, r = 3399) and
// thrust::uniform_real_distribution<float>(0,1): float(x - 1) * 2^-31.
__device__ __forceinline__ float rng_u01(uint32_t& x)
{
    const uint32_t q = x / 44488u;
    const uint32_t t1 = (x - q * 44488u) * 48271u;
    const uint32_t t2 = q * 3399u;
    x = t1 < t2 ? t1 + (2147483647u - t2) : t1 - t2;
    return __fmul_rn(__uint2float_rn(x - 1u), 4.656612873077392578125e-10f);
}

// ---- scene records ---------------------------------------------------------------------------
// Compact primitive record (160 B, 16-B aligned) built on the host from struct Geom:
// column c, row r of a glm::mat4 m is m[c*4 + r]; only rows 0..2 are ever used.
struct PrimRec {
    float inv[12];   // inverseTransform  [c*3 + r]
    float fwd[12];   // transform
    float itr[12];   // invTranspose
    int type;
    int material;
    int pad0, pad1;
};
static_assert(sizeof(PrimRec) == 160, "PrimRec must stay 160 bytes");

struct MatRec {      // == struct Material, 48 B
    float color[3];
    float spec_exp;
    float spec_color[3];
    float has_reflective;
    float has_refractive;
    float ior;
    float emittance;
    int texture_id;
};
static_assert(sizeof(MatRec) == 48, "MatRec must stay 48 bytes");

// glm::mat4 * vec4(v, 1) -> xyz  (operator*, GLM 0.9.6 type_mat4x4.inl; after contraction:
// add(fma(x, m0, y*m1), fma(z, m2, m3)))
__device__ __forceinline__ V3 xform_point(const float* __restrict__ m, const V3& v)
{
    return v3(__fadd_rn(__fmaf_rn(v.x, m[0], __fmul_rn(v.y, m[3])), __fmaf_rn(v.z, m[6], m[9])),
              __fadd_rn(__fmaf_rn(v.x, m[1], __fmul_rn(v.y, m[4])), __fmaf_rn(v.z, m[7], m[10])),
              __fadd_rn(__fmaf_rn(v.x, m[2], __fmul_rn(v.y, m[5])), __fmaf_rn(v.z, m[8], m[11])));
}

// glm::mat4 * vec4(v, 0) -> xyz: add(fma(x, m0, y*m1), fma(z, m2, m3*0))
__device__ __forceinline__ V3 xform_vector(const float* __restrict__ m, const V3& v)
{
    return v3(__fadd_rn(__fmaf_rn(v.x, m[0], __fmul_rn(v.y, m[3])), __fmaf_rn(v.z, m[6], __fmul_rn(m[9], 0.0f))),
              __fadd_rn(__fmaf_rn(v.x, m[1], __fmul_rn(v.y, m[4])), __fmaf_rn(v.z, m[7], __fmul_rn(m[10], 0.0f))),
              __fadd_rn(__fmaf_rn(v.x, m[2], __fmul_rn(v.y, m[5])), __fmaf_rn(v.z, m[8], __fmul_rn(m[11], 0.0f))));
}

// getPointOnRay (intersections.h:29-32): o + (t - 1e-4) * normalize(d), fused as fma
__device__ __forceinline__ V3 point_on_ray(const V3& o, const V3& d, float t)
{
    const float tt = __fadd_rn(t, -1.0e-4f);
    const V3 nd = normalize_exact(d);
    return v3(__fmaf_rn(tt, nd.x, o.x), __fmaf_rn(tt, nd.y, o.y), __fmaf_rn(tt, nd.z, o.z));
}

// Candidate hit of one primitive: world distance t (<= 0 / NaN = no hit) and what is needed to
// build the normal later (only the closest hit's normal is ever used).
struct Cand {
    float t;
    V3 aux;        // cube: the slab normal in object space; sphere: the object-space hit point
    bool outside;
};

// boxIntersectionTest (intersections.cu:4-58) on an object-space ray (qo, qd normalised).
__device__ __forceinline__ bool box_test(const V3& qo, const V3& qd, float& t_obj, V3& n_obj)
{
    float tmin = -1e38f, tmax = 1e38f;
    V3 tmin_n = v3(0.f, 0.f, 0.f), tmax_n = v3(0.f, 0.f, 0.f);
    const float o[3] = {qo.x, qo.y, qo.z};
    const float d[3] = {qd.x, qd.y, qd.z};
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float t1 = __fdiv_rn(__fsub_rn(-0.5f, o[a]), d[a]);
        const float t2 = __fdiv_rn(__fsub_rn(0.5f, o[a]), d[a]);
        const float ta = t1 < t2 ? t1 : t2;      // glm::min
        const float tb = t1 > t2 ? t1 : t2;      // glm::max
        const float s = t2 < t1 ? 1.0f : -1.0f;
        const V3 n = v3(a == 0 ? s : 0.f, a == 1 ? s : 0.f, a == 2 ? s : 0.f);
        if (ta > 0.f && ta > tmin) { tmin = ta; tmin_n = n; }
        if (tb < tmax) { tmax = tb; tmax_n = n; }
    }
    if (tmax >= tmin && tmax > 0.f) {
        if (tmin <= 0.f) { tmin = tmax; tmin_n = tmax_n; }
        t_obj = tmin;
        n_obj = tmin_n;
        return true;
    }
    return false;
}

// sphereIntersectionTest (intersections.cu:60-115) on an object-space ray.
__device__ __forceinline__ bool sphere_test(const V3& ro, const V3& rd, float& t_obj, bool& outside)
{
    const float vdd = dot_exact(ro, rd);
    const float doo = dot_exact(ro, ro);
    const float radicand = __fmaf_rn(vdd, vdd, __fsub_rn(0.25f, doo));   // powf(.5, 2) == 0.25
    if (radicand < 0.f) return false;
    const float sq = __fsqrt_rn(radicand);
    const float t1 = __fsub_rn(sq, vdd);
    const float t2 = __fsub_rn(-vdd, sq);
    if (t1 < 0.f && t2 < 0.f) return false;
    if (t1 > 0.f && t2 > 0.f) { t_obj = fminf(t1, t2); outside = true; }
    else { t_obj = fmaxf(t1, t2); outside = false; }
    return true;
}

// One primitive against a world-space ray: what computeIntersections' loop body does through
// box/sphereIntersectionTest (pathtrace.cu:252-258), minus the normal.
__device__ __forceinline__ Cand prim_test(const PrimRec& g, const V3& o, const V3& d)
{
    Cand c;
    c.t = -1.0f;
    c.outside = true;
    c.aux = v3(0.f, 0.f, 0.f);
    const V3 qo = xform_point(g.inv, o);
    const V3 qd = normalize_exact(xform_vector(g.inv, d));
    float t_obj;
    if (g.type == 1) {                 // CUBE
        V3 n;
        if (!box_test(qo, qd, t_obj, n)) return c;
        c.aux = n;
        const V3 w = xform_point(g.fwd, point_on_ray(qo, qd, t_obj));
        c.t = distance_exact(o, w);
    } else if (g.type == 0) {          // SPHERE
        bool outside;
        if (!sphere_test(qo, qd, t_obj, outside)) return c;
        const V3 p = point_on_ray(qo, qd, t_obj);
        c.aux = p;
        c.outside = outside;
        const V3 w = xform_point(g.fwd, p);
        c.t = distance_exact(o, w);
    }
    return c;
}

// normal = normalize(invTranspose * vec4(aux, 0)); the sphere flips it when hit from inside.
__device__ __forceinline__ V3 prim_normal(const PrimRec& g, const Cand& c)
{
    V3 n = normalize_exact(xform_vector(g.itr, c.aux));
    if (g.type == 0 && !c.outside) n = v3(-n.x, -n.y, -n.z);
    return n;
}

}  // namespace ptk
