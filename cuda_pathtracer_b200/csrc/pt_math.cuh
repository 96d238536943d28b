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
struct alignas(16) PrimRec {
    float inv[12];   // inverseTransform  [c*3 + r]
    float fwd[12];   // transform
    float itr[12];   // invTranspose
    int type;
    int material;
    int pad0, pad1;
};
static_assert(sizeof(PrimRec) == 160, "PrimRec must stay 160 bytes");

struct alignas(16) MatRec {      // == struct Material, 48 B
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
// GLOBAL: the record is in global memory -- its matrices are then taken with 16-byte ld.global.nc loads (three for the
// inverse, one for the type word, three more for the forward matrix once the ray has hit) instead of two dozen 4-byte ones:
// in the tree walks every lane tests another primitive, and each load instruction is 32 tag look-ups.
template <bool GLOBAL>
__device__ __forceinline__ Cand prim_test_t(const PrimRec& g, const V3& o, const V3& d)
{
    Cand c;
    c.t = -1.0f;
    c.outside = true;
    c.aux = v3(0.f, 0.f, 0.f);
    float inv[12];
    int type;
    const float4* q = reinterpret_cast<const float4*>(&g);
    if (GLOBAL) {
        const float4 m0 = __ldg(q), m1 = __ldg(q + 1), m2 = __ldg(q + 2), w = __ldg(q + 9);
        inv[0] = m0.x; inv[1] = m0.y; inv[2] = m0.z; inv[3] = m0.w; inv[4] = m1.x; inv[5] = m1.y; inv[6] = m1.z; inv[7] = m1.w;
        inv[8] = m2.x; inv[9] = m2.y; inv[10] = m2.z; inv[11] = m2.w;
        type = __float_as_int(w.x);
    } else {
#pragma unroll
        for (int k = 0; k < 12; ++k) inv[k] = g.inv[k];
        type = g.type;
    }
    const V3 qo = xform_point(inv, o);
    const V3 qd = normalize_exact(xform_vector(inv, d));
    float t_obj;
    V3 p;
    if (type == 1) {                   // CUBE
        V3 n;
        if (!box_test(qo, qd, t_obj, n)) return c;
        c.aux = n;
        p = point_on_ray(qo, qd, t_obj);
    } else if (type == 0) {            // SPHERE
        bool outside;
        if (!sphere_test(qo, qd, t_obj, outside)) return c;
        p = point_on_ray(qo, qd, t_obj);
        c.aux = p;
        c.outside = outside;
    } else {
        return c;
    }
    float fwd[12];
    if (GLOBAL) {
        const float4 m0 = __ldg(q + 3), m1 = __ldg(q + 4), m2 = __ldg(q + 5);
        fwd[0] = m0.x; fwd[1] = m0.y; fwd[2] = m0.z; fwd[3] = m0.w; fwd[4] = m1.x; fwd[5] = m1.y; fwd[6] = m1.z; fwd[7] = m1.w;
        fwd[8] = m2.x; fwd[9] = m2.y; fwd[10] = m2.z; fwd[11] = m2.w;
    } else {
#pragma unroll
        for (int k = 0; k < 12; ++k) fwd[k] = g.fwd[k];
    }
    c.t = distance_exact(o, xform_point(fwd, p));
    return c;
}
__device__ __forceinline__ Cand prim_test(const PrimRec& g, const V3& o, const V3& d) { return prim_test_t<false>(g, o, d); }

// normal = normalize(invTranspose * vec4(aux, 0)); the sphere flips it when hit from inside.
__device__ __forceinline__ V3 prim_normal(const float* __restrict__ itr, int type, const Cand& c)
{
    V3 n = normalize_exact(xform_vector(itr, c.aux));
    if (type == 0 && !c.outside) n = v3(-n.x, -n.y, -n.z);
    return n;
}
__device__ __forceinline__ V3 prim_normal(const PrimRec& g, const Cand& c) { return prim_normal(g.itr, g.type, c); }

// The last 64 bytes of a PrimRec -- invTranspose, type, material: all that is read of the WINNING primitive -- taken into
// registers through an address-space specific path (four 16-byte loads): the winner's record is in shared memory for a
// staged primitive and in global memory otherwise, and a reference bound to `staged ? sprims[k] : prims[hit]` is a generic
// pointer (a dozen 4-byte generic loads behind ten instructions of address conversion, see lds_f4 below).
struct PrimTail {
    float itr[12];
    int type, material, pad0, pad1;
};


// ---- triangle meshes: BVH traversal and ray / triangle test ---------------------------------------
// BVHIntersectionTest (intersections.cu:170-224), BoundingBox::intersect (boundingbox.h:74-92),
// Triangle::intersect (sceneStructs.h:146-160) and glm::intersectRayTriangle (GLM 0.9.6
// gtx/intersect.inl), in the operation order of the reference build's SASS
// (_Z19BVHIntersectionTest...): cross products as FMUL + FFMA(a, b, -t), dot products as
// fma(z, z', fma(x, x', y*y')), 1/x as rcp.rn, every bounding-box slab as FADD then FMUL.

// Address-space specific 16-byte loads for the tree walks.  A node comes from the staged top of its tree (shared memory) or
// from the array in global memory; written as `cond ? snodes[i] : gnodes[i]` the compiler selects between two GENERIC
// addresses -- a divergent branch, eleven instructions to turn the shared offset into a generic address (S2R CgaCtaId, S2R
// SWINHI, ...) on one side, and generic loads behind it: 14-17 issue slots per node visit, 5-7 % of the walk kernels
// (profiles/r2z_config5_k_isect_dyn_lines.txt: pt_kernels.cu:1178).  With these a side is one address add and its loads.
__device__ __forceinline__ float4 lds_f4(unsigned addr)
{
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint4 lds_u4(unsigned addr)
{
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}

__device__ __forceinline__ PrimTail load_prim_tail(bool in_shared, unsigned smem_addr, const PrimRec* __restrict__ g)
{
    float4 w[4];
    if (in_shared) {
#pragma unroll
        for (int j = 0; j < 4; ++j) w[j] = lds_f4(smem_addr + 96u + 16u * j);
    } else {
        const float4* q = reinterpret_cast<const float4*>(g) + 6;
#pragma unroll
        for (int j = 0; j < 4; ++j) w[j] = __ldg(q + j);
    }
    PrimTail t;
    t.itr[0] = w[0].x; t.itr[1] = w[0].y; t.itr[2] = w[0].z; t.itr[3] = w[0].w;
    t.itr[4] = w[1].x; t.itr[5] = w[1].y; t.itr[6] = w[1].z; t.itr[7] = w[1].w;
    t.itr[8] = w[2].x; t.itr[9] = w[2].y; t.itr[10] = w[2].z; t.itr[11] = w[2].w;
    t.type = __float_as_int(w[3].x); t.material = __float_as_int(w[3].y);
    t.pad0 = __float_as_int(w[3].z); t.pad1 = __float_as_int(w[3].w);
    return t;
}

// a material (48 bytes) the same way: three 16-byte loads from the staged copy or from global memory
__device__ __forceinline__ MatRec load_mat(bool in_shared, unsigned smem_addr, const MatRec* __restrict__ g)
{
    float4 w[3];
    if (in_shared) {
#pragma unroll
        for (int j = 0; j < 3; ++j) w[j] = lds_f4(smem_addr + 16u * j);
    } else {
        const float4* q = reinterpret_cast<const float4*>(g);
#pragma unroll
        for (int j = 0; j < 3; ++j) w[j] = __ldg(q + j);
    }
    MatRec m;
    m.color[0] = w[0].x; m.color[1] = w[0].y; m.color[2] = w[0].z; m.spec_exp = w[0].w;
    m.spec_color[0] = w[1].x; m.spec_color[1] = w[1].y; m.spec_color[2] = w[1].z; m.has_reflective = w[1].w;
    m.has_refractive = w[2].x; m.ior = w[2].y; m.emittance = w[2].z; m.texture_id = __float_as_int(w[2].w);
    return m;
}

struct alignas(16) BvhRec {          // class BVHTree (40 B) padded to three 16-byte words
    float lo[3]; int sub_areas;          // min corner | leaf triangle count (0 = inner node)
    float hi[3]; int axis;               // max corner | split axis of an inner node
    int first; int rchild; int pad0, pad1;
};
static_assert(sizeof(BvhRec) == 48, "BvhRec must stay 48 bytes");

struct alignas(16) TriVert {         // what the intersection test reads: 48 B
    float v0[3]; int id;                 // id = index before the BVH build reordered the triangles
    float v1[3]; int pad0;
    float v2[3]; int pad1;
};
static_assert(sizeof(TriVert) == 48, "TriVert must stay 48 bytes");

struct alignas(16) TriAttr {         // what only the closest hit reads: 64 B
    float uv[3][2];
    float n[3][3];
    float pad;
};
static_assert(sizeof(TriAttr) == 64, "TriAttr must stay 64 bytes");

struct TexRec {          // class Texture: size, channel count, byte offset into the texel pool
    int w, h, comp;
    unsigned offset;
};

struct MeshScene {
    const BvhRec* nodes;
    const TriVert* tv;
    const TriAttr* ta;
    const TexRec* tex;
    const unsigned char* texels;
    int n_nodes, n_tris, n_tex, n_smem_nodes;
    int literal_walk;    // the tree is >= BVH_MAX_SEARCH_DEPTH levels deep: keep the reference's stack-overflow behaviour
    int brute;           // GuiDataContainer::useBVHtree == false: 1 = meshIntersectionTest with its bounding-box cull
                         // (useBBox), 2 = without; 0 = the BVH walk
    int pad_[2];
};

struct MeshHit {
    bool any;            // some triangle was intersected (what BVHIntersectionTest returns)
    float t;             // parametric distance of the closest one (ShadeableIntersection::t)
    int tri;             // its position in the reordered triangle array (-1: none recorded)
    int id;              // its Triangle::id
    float bx, by;        // barycentric coordinates of the hit
};

// glm::min / glm::max: x < y ? x : y  /  x > y ? x : y (func_common.inl:413,434)
__device__ __forceinline__ float gmin(float x, float y) { return x < y ? x : y; }
__device__ __forceinline__ float gmax(float x, float y) { return x > y ? x : y; }

// BoundingBox::intersect (boundingbox.h:74-92).  `entry` (only meaningful on a hit) is the slab
// entry distance the test computes anyway.
__device__ __forceinline__ bool bbox_hit(const float* __restrict__ lo, const float* __restrict__ hi, const V3& o,
                                         const V3& inv, float& entry)
{
    const float mx = __fmul_rn(__fsub_rn(lo[0], o.x), inv.x), Mx = __fmul_rn(__fsub_rn(hi[0], o.x), inv.x);
    const float my = __fmul_rn(__fsub_rn(lo[1], o.y), inv.y), My = __fmul_rn(__fsub_rn(hi[1], o.y), inv.y);
    const float mz = __fmul_rn(__fsub_rn(lo[2], o.z), inv.z), Mz = __fmul_rn(__fsub_rn(hi[2], o.z), inv.z);
    const float tmin = gmax(gmax(gmin(mx, Mx), gmin(my, My)), gmin(mz, Mz));
    const float tmax = gmin(gmin(gmax(mx, Mx), gmax(my, My)), gmax(mz, Mz));
    entry = tmin;
    if (tmax < 0.0f) return false;
    if (tmin > tmax) return false;
    return true;
}

// glm::intersectRayTriangle: back-face culled (a < epsilon -> miss); bary = (x, y, t)
__device__ __forceinline__ bool tri_test(const TriVert& T, const V3& o, const V3& d, float& bx, float& by, float& bt)
{
    const V3 v0 = v3(T.v0[0], T.v0[1], T.v0[2]);
    const V3 e1 = v3(__fsub_rn(T.v1[0], v0.x), __fsub_rn(T.v1[1], v0.y), __fsub_rn(T.v1[2], v0.z));
    const V3 e2 = v3(__fsub_rn(T.v2[0], v0.x), __fsub_rn(T.v2[1], v0.y), __fsub_rn(T.v2[2], v0.z));
    const V3 p = cross_exact(d, e2);
    const float a = dot_exact(e1, p);
    if (a < 1.1920928955078125e-07f) return false;
    const float f = __frcp_rn(a);
    const V3 s = v3(__fsub_rn(o.x, v0.x), __fsub_rn(o.y, v0.y), __fsub_rn(o.z, v0.z));
    bx = __fmul_rn(dot_exact(s, p), f);
    if (bx < 0.0f) return false;
    if (bx > 1.0f) return false;
    const V3 q = cross_exact(s, e1);
    by = __fmul_rn(dot_exact(d, q), f);
    if (by < 0.0f) return false;
    if (__fadd_rn(bx, by) > 1.0f) return false;
    bt = __fmul_rn(dot_exact(e2, q), f);
    return bt >= 0.0f;
}

// the 48 bytes of a triangle's vertices as three 16-byte loads (nine 4-byte loads otherwise: in a diverged walk every lane is at
// another triangle, and each load instruction is 32 tag look-ups)
__device__ __forceinline__ TriVert load_tri(const TriVert* __restrict__ tv, int at)
{
    const float4* q = reinterpret_cast<const float4*>(tv + at);
    const float4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
    TriVert T;
    T.v0[0] = a.x; T.v0[1] = a.y; T.v0[2] = a.z; T.id = __float_as_int(a.w);
    T.v1[0] = b.x; T.v1[1] = b.y; T.v1[2] = b.z; T.pad0 = 0;
    T.v2[0] = c.x; T.v2[1] = c.y; T.v2[2] = c.z; T.pad1 = 0;
    return T;
}

constexpr int BVH_MAX_SEARCH_DEPTH = 64;     // intersections.cu:177

// One traversal of the BVH over ALL meshes' triangles.  The reference runs this once per MESH geom
// and accepts the result only for the geom whose id range holds the hit triangle
// (pathtrace.cu:262-272); the traversal itself does not depend on the geom, so it is done once.
// snodes: the first M.n_smem_nodes nodes staged in shared memory (DFS order keeps the top of the
// tree there).
__device__ __forceinline__ void tri_candidate(const MeshScene& M, int at, const V3& o, const V3& d, MeshHit& h)
{
    float bx, by, bt;
    if (tri_test(load_tri(M.tv, at), o, d, bx, by, bt)) {
        h.any = true;
        if (h.t == -1.0f || bt < h.t) {      // strict: the first one met wins a tie
            h.t = bt; h.tri = at; h.bx = bx; h.by = by;
        }
    }
}

// The reference's walk, literally: leaves are tested as they are met, no pruning by distance, and a
// full stack silently drops the node at hand (intersections.cu:203-206).  Used for trees deep
// enough (>= BVH_MAX_SEARCH_DEPTH levels) for that drop to be reachable.
__device__ __forceinline__ MeshHit bvh_traverse_literal(const MeshScene& M, const BvhRec* __restrict__ snodes, const V3& o,
                                                        const V3& d)
{
    MeshHit h;
    h.any = false; h.t = 3.402823466e+38f; h.tri = -1; h.id = -1; h.bx = 0.f; h.by = 0.f;
    int to_visit[BVH_MAX_SEARCH_DEPTH];
    int sp = 0, cur = 0;
    const bool neg[3] = {d.x < 0.0f, d.y < 0.0f, d.z < 0.0f};
    const V3 inv = v3(__frcp_rn(d.x), __frcp_rn(d.y), __frcp_rn(d.z));
    while (true) {
        const BvhRec& nd = cur < M.n_smem_nodes ? snodes[cur] : M.nodes[cur];
        float entry;
        if (bbox_hit(nd.lo, nd.hi, o, inv, entry)) {
            const int n_leaf = nd.sub_areas;
            if (n_leaf > 0) {
                for (int i = 0; i < n_leaf; ++i) tri_candidate(M, nd.first + i, o, d, h);
                if (sp == 0) break;
                cur = to_visit[--sp];
            } else {
                if (sp == BVH_MAX_SEARCH_DEPTH) { cur = to_visit[--sp]; continue; }
                if (neg[nd.axis]) { to_visit[sp++] = cur + 1; cur = nd.rchild; }
                else { to_visit[sp++] = nd.rchild; cur = cur + 1; }
            }
        } else {
            if (sp == 0) break;
            cur = to_visit[--sp];
        }
    }
    if (h.tri >= 0) h.id = M.tv[h.tri].id;
    return h;
}

// Same result, arranged for the SIMT machine.  The reference's walk visits the same nodes in the
// same order whatever it has hit so far, and keeps the nearest triangle, the first one met on a
// tie.  So (1) the triangle tests of the leaves can be deferred: the walk only runs the (exact)
// box tests and lists the leaves it reaches, then the listed triangles are tested in the order
// met -- one triangle per lane per trip, instead of a whole leaf whenever ANY lane of the warp
// stands at one; (2) a node or leaf whose box starts behind the current hit (with a margin far
// above the rounding of the slab distances) cannot hold a nearer or equal one and is skipped.
#ifndef PT_MLEAVES_N
#define PT_MLEAVES_N 12
#endif
constexpr int PT_MLEAVES = PT_MLEAVES_N;      // listed leaves before the list is worked off

// a distance safely beyond t (t may be negative: glm::intersectRayTriangle also reports hits behind the origin)
__device__ __forceinline__ float prune_limit(float t) { return fmaf(fabsf(t), 1e-4f, t) + 1e-4f; }

// t_cap: nothing at or beyond this distance is of interest to the caller (+inf: everything is).
__device__ __forceinline__ MeshHit bvh_traverse(const MeshScene& M, const BvhRec* __restrict__ snodes, const V3& o,
                                                const V3& d, float t_cap = 3.402823466e+38f)
{
    MeshHit h;
    h.any = false; h.t = 3.402823466e+38f; h.tri = -1; h.id = -1; h.bx = 0.f; h.by = 0.f;
    if (M.n_nodes == 0) return h;
    if (M.literal_walk) return bvh_traverse_literal(M, snodes, o, d);
    int to_visit[BVH_MAX_SEARCH_DEPTH];
    int leaf_first[PT_MLEAVES], leaf_count[PT_MLEAVES];
    float leaf_entry[PT_MLEAVES];
    int sp = 0, cur = 0, nl = 0;
    const bool neg[3] = {d.x < 0.0f, d.y < 0.0f, d.z < 0.0f};
    const V3 inv = v3(__frcp_rn(d.x), __frcp_rn(d.y), __frcp_rn(d.z));
    bool more = true;
    const unsigned snodes_s = (unsigned)__cvta_generic_to_shared(snodes);
    while (more) {
        const float lim = fminf(prune_limit(h.t), t_cap);
        while (true) {                                   // the walk: box tests only
            // (two address-space specific paths instead of a select between generic addresses: see lds_f4)
            float4 w0, w1, w2;                           // BvhRec as three 16-byte words
            if (cur < M.n_smem_nodes) {
                const unsigned at = snodes_s + 48u * (unsigned)cur;
                w0 = lds_f4(at); w1 = lds_f4(at + 16u); w2 = lds_f4(at + 32u);
            } else {
                const float4* g = reinterpret_cast<const float4*>(M.nodes + cur);
                w0 = __ldg(g); w1 = __ldg(g + 1); w2 = __ldg(g + 2);
            }
            const float nlo[3] = {w0.x, w0.y, w0.z}, nhi[3] = {w1.x, w1.y, w1.z};
            const int nd_sub = __float_as_int(w0.w), nd_axis = __float_as_int(w1.w);
            const int nd_first = __float_as_int(w2.x), nd_rchild = __float_as_int(w2.y);
            float entry;
            bool descend = false;
            if (bbox_hit(nlo, nhi, o, inv, entry) && !(entry > lim)) {
                if (nd_sub > 0) {
                    leaf_first[nl] = nd_first; leaf_count[nl] = nd_sub; leaf_entry[nl] = entry; ++nl;
                } else {
                    descend = true;
                    if (neg[nd_axis]) { to_visit[sp++] = cur + 1; cur = nd_rchild; }
                    else { to_visit[sp++] = nd_rchild; cur = cur + 1; }
                }
            }
            if (!descend) {
                if (sp == 0) { more = false; break; }
                cur = to_visit[--sp];
            }
            if (nl == PT_MLEAVES) break;
        }
        int k = 0, i = 0;                                // the listed triangles, in the order met
        while (true) {
            const float lim2 = fminf(prune_limit(h.t), t_cap);
            while (k < nl && (i >= leaf_count[k] || leaf_entry[k] > lim2)) { ++k; i = 0; }
            if (k >= nl) break;
            tri_candidate(M, leaf_first[k] + i, o, d, h);
            ++i;
        }
        nl = 0;
    }
    if (h.tri >= 0) h.id = M.tv[h.tri].id;
    return h;
}

// meshIntersectionTest (intersections.cu:120-167): what computeIntersections calls for a MESH primitive when
// settings->use_BVHtree is false (pathtrace.cu:273-275).  Optional slab test against the primitive's world
// bounds (kept in the otherwise unused `fwd` words of a MESH PrimRec), then EVERY triangle at positions
// [pad0, pad1) of the (BVH-reordered, sic) triangle array: nearest positive distance, first one on a tie.
// Arithmetic order from the reference build's SASS (_Z20meshIntersectionTest...): rcp.rn for 1/d, the slabs
// as FADD then FMUL, the same ray / triangle code as the BVH path.
__device__ __forceinline__ bool mesh_brute(const MeshScene& M, const PrimRec& g, const V3& o, const V3& d, MeshHit& h)
{
    if (M.brute == 1) {
        const V3 inv = v3(__frcp_rn(d.x), __frcp_rn(d.y), __frcp_rn(d.z));
        float entry;
        if (!bbox_hit(g.fwd, g.fwd + 3, o, inv, entry)) return false;
    }
    h.any = false; h.t = 3.402823466e+38f; h.tri = -1; h.id = -1; h.bx = 0.f; h.by = 0.f;
    const int end = g.pad1 < M.n_tris ? g.pad1 : M.n_tris;
    for (int at = g.pad0; at < end; ++at) {
        float bx, by, bt;
        if (tri_test(load_tri(M.tv, at), o, d, bx, by, bt) && bt > 0.0f && bt < h.t) {
            h.any = true; h.t = bt; h.tri = at; h.bx = bx; h.by = by;
        }
    }
    return h.any;
}

// its attribute blend (sic, intersections.cu:160-164): actual_z * a0 + a1 + a2 -- not a barycentric one
__device__ __forceinline__ void brute_attributes(const TriAttr& A, float bx, float by, V3& n, float& u, float& v)
{
    const float w = __fsub_rn(__fsub_rn(1.0f, bx), by);
    const V3 m = v3(__fadd_rn(__fmaf_rn(w, A.n[0][0], A.n[1][0]), A.n[2][0]),
                    __fadd_rn(__fmaf_rn(w, A.n[0][1], A.n[1][1]), A.n[2][1]),
                    __fadd_rn(__fmaf_rn(w, A.n[0][2], A.n[1][2]), A.n[2][2]));
    n = normalize_exact(m);
    u = __fadd_rn(__fmaf_rn(w, A.uv[0][0], A.uv[1][0]), A.uv[2][0]);
    v = __fadd_rn(__fmaf_rn(w, A.uv[0][1], A.uv[1][1]), A.uv[2][1]);
}

// Triangle::intersect's attribute interpolation for the closest hit (sceneStructs.h:152-154):
// uv = uv0*(1-x-y) + uv1*x + uv2*y; normal = normalize(n0*x + n1*y + n2*(1-x-y)) -- the weights of
// the normal are permuted with respect to the uv's (sic).
// the 64 bytes of a triangle's attributes as four 16-byte loads
__device__ __forceinline__ TriAttr load_attr(const TriAttr* __restrict__ ta, int at)
{
    const float4* q = reinterpret_cast<const float4*>(ta + at);
    const float4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2), d = __ldg(q + 3);
    TriAttr A;
    A.uv[0][0] = a.x; A.uv[0][1] = a.y; A.uv[1][0] = a.z; A.uv[1][1] = a.w;
    A.uv[2][0] = b.x; A.uv[2][1] = b.y; A.n[0][0] = b.z; A.n[0][1] = b.w;
    A.n[0][2] = c.x; A.n[1][0] = c.y; A.n[1][1] = c.z; A.n[1][2] = c.w;
    A.n[2][0] = d.x; A.n[2][1] = d.y; A.n[2][2] = d.z; A.pad = d.w;
    return A;
}

__device__ __forceinline__ void tri_attributes(const TriAttr& A, float bx, float by, V3& n, float& u, float& v)
{
    const float w = __fsub_rn(__fsub_rn(1.0f, bx), by);
    u = __fmaf_rn(by, A.uv[2][0], __fmaf_rn(w, A.uv[0][0], __fmul_rn(bx, A.uv[1][0])));
    v = __fmaf_rn(by, A.uv[2][1], __fmaf_rn(w, A.uv[0][1], __fmul_rn(bx, A.uv[1][1])));
    const V3 m = v3(__fmaf_rn(w, A.n[2][0], __fmaf_rn(bx, A.n[0][0], __fmul_rn(by, A.n[1][0]))),
                    __fmaf_rn(w, A.n[2][1], __fmaf_rn(bx, A.n[0][1], __fmul_rn(by, A.n[1][1]))),
                    __fmaf_rn(w, A.n[2][2], __fmaf_rn(bx, A.n[0][2], __fmul_rn(by, A.n[1][2]))));
    n = normalize_exact(m);
}

// Texture::get_color (sceneStructs.h:175-188): nearest texel, v flipped; only 3-channel textures
// yield a colour.  (The reference reads out of bounds for uv outside [0, 1]; the index is clamped here.)
__device__ __forceinline__ V3 tex_color(const MeshScene& M, int texture_id, float u, float v)
{
    const TexRec T = M.tex[texture_id];
    const float fw = (float)T.w, fh = (float)T.h;
    const int X = (int)gmin(__fmul_rn(fw, u), __fadd_rn(fw, -1.0f));
    const int Y = (int)gmin(__fmul_rn(fh, __fsub_rn(1.0f, v)), __fadd_rn(fh, -1.0f));
    if (T.comp != 3) return v3(0.f, 0.f, 0.f);
    int texid = Y * T.w + X;
    texid = texid < 0 ? 0 : (texid >= T.w * T.h ? T.w * T.h - 1 : texid);
    const unsigned char* px = M.texels + T.offset + 3u * (unsigned)texid;
    return v3(__fmul_rn((float)px[0], 0.003921568627f), __fmul_rn((float)px[1], 0.003921568627f),
              __fmul_rn((float)px[2], 0.003921568627f));
}

}  // namespace ptk
