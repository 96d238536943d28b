// pt_kernels.cu -- hand-written sm_100a kernels of the wavefront path tracer.
//
// Replaces generateRayFromCamera, computeIntersections, shadeMaterials,
// mark_valid + Efficient::scan + keep, finalGather and sendImageToPBO of
// path_tracer/src/pathtrace.cu with:
//
//   k_raygen   one camera path per pixel into the SoA pool (three 16-B stores per path)
//   k_bounce   ONE kernel per bounce for every iteration in flight: closest hit over the
//              primitive records (staged into shared memory with TMA bulk copies,
//              cp.async.bulk + mbarrier, double-buffered when the scene has more than one
//              chunk), shading / BSDF sampling / Russian roulette in registers, the terminated
//              paths' colour accumulated straight into the image, and the survivors written
//              to the other pool through a single-pass decoupled look-back STABLE compaction
//              (warp ballot ranks, one 64-bit descriptor per tile).  The hit record never
//              touches HBM and the live count stays on the device (no host round trip).
//   k_tonemap  sendImageToPBO into a headless uchar4 buffer.
//
// Roofline: the Cornell-size scenes are FP32-issue bound (about 1.8 k instructions per path
// segment: IEEE div/sqrt/rcp sequences that exact parity requires) with HBM close behind;
// algorithmic bytes are 96 B per path segment + 24 B per pixel per iteration.
#include <cstdio>
#include "pt_kernels.cuh"
#include "lookback.cuh"

namespace ptk {

using b200::FULL_MASK;

namespace {

__device__ __forceinline__ unsigned smem_u32(const void* p)
{
    return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
// TMA 1-D bulk copy global -> shared, completion signalled on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
#ifndef PT_MBAR_HINT_NS
#define PT_MBAR_HINT_NS 0
#endif
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity)
{
#if PT_MBAR_HINT_NS > 0
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity), "r"((unsigned)PT_MBAR_HINT_NS)
        : "memory");
#else
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
#endif
}

// The small-scene search's records (closest_small) live in STATIC shared memory: their addresses are compile-time constants,
// so the cull loop's loads are LDS [k * 32 + constant].  Behind the run-time sized arrays of the dynamic allocation (materials,
// mesh nodes) ptxas rebuilt the base address -- ten instructions -- for every primitive of every ray, 7 % of k_bounce's
// instructions on the Cornell box (profiles/r2z_config2_cornell_k_bounce_src.csv).  cull == prims + PT_SMALL_MAX.
template <bool ON>
struct alignas(16) SmallStageT {
    PrimRec prims[PT_SMALL_MAX];
    float4 cull[2 * PT_SMALL_MAX];
};
template <>
struct alignas(16) SmallStageT<false> {
    PrimRec prims[1];
};

struct BounceShared {
    unsigned long long bar[2];
    int warp_cnt[PT_THREADS / 32];
    int tile_exclusive;
    int pad[3];
};

// closest hit over the primitives [0, n) of `sp`, global indices base..base+n-1
// A MESH primitive's candidate is the BVH hit if the hit triangle belongs to it (pathtrace.cu:262-272).
// `owner` records the first MESH primitive that accepted the hit.
// M.brute (useBVHtree == false, pathtrace.cu:273-275): every MESH primitive runs meshIntersectionTest instead;
// `last` is what tmp_normal / tmp_coord hold (written only by a mesh test that hits) and `mh` what the current
// winner copied from them -- so a sphere / cube winner carries the texture coordinate of the LAST mesh
// primitive before it that was hit.
__device__ __forceinline__ void closest_in_chunk(const PrimRec* __restrict__ sp, int n, int base, const V3& o,
                                                 const V3& d, const MeshScene& M, MeshHit& mh, MeshHit& last, float& t_min,
                                                 int& hit, Cand& best, int& owner)
{
    for (int i = 0; i < n; ++i) {
        Cand c = prim_test(sp[i], o, d);
        if (sp[i].type == 2) {
            if (M.brute) {
                MeshHit h;
                if (mesh_brute(M, sp[i], o, d, h)) { last = h; c.t = h.t; }
            } else if (mh.any && mh.id >= sp[i].pad0 && mh.id < sp[i].pad1) {
                c.t = mh.t;
                if (owner < 0) owner = base + i;
            }
        }
        if (c.t > 0.0f && t_min > c.t) {    // strict: the lowest index wins ties (pathtrace.cu:280)
            t_min = c.t;
            hit = base + i;
            best = c;
            if (M.brute) mh = last;
        }
    }
}

// normal / texture coordinate of the winner `g` (scene index `hit`), pathtrace.cu:283-296
__device__ __forceinline__ void hit_attributes(const MeshScene& M, const PrimTail& g, int hit, const Cand& best, const MeshHit& mh,
                                               int owner, V3& n, float& u, float& v)
{
    if (M.brute) {
        if (g.type == 2) {
            brute_attributes(M.ta[mh.tri], mh.bx, mh.by, n, u, v);
        } else {
            n = prim_normal(g.itr, g.type, best);
            if (mh.tri >= 0) { V3 unused; brute_attributes(M.ta[mh.tri], mh.bx, mh.by, unused, u, v); }
        }
    } else if (g.type == 2) {
        tri_attributes(load_attr(M.ta, mh.tri), mh.bx, mh.by, n, u, v);
    } else {
        n = prim_normal(g.itr, g.type, best);
        // (sic) a sphere / cube winning AFTER a mesh primitive accepted the BVH hit carries that
        // hit's stale texture coordinate (pathtrace.cu:267,283)
        if (owner >= 0 && owner < hit) { V3 unused; tri_attributes(load_attr(M.ta, mh.tri), mh.bx, mh.by, unused, u, v); }
    }
}

// one 16-bit grid coordinate of a 32-byte geometry-BVH node -> float: 0x4B000000 | q is the float 2^23 + q exactly, so the
// integer-to-float conversion (a quarter-rate instruction) becomes a byte permute and an add
__device__ __forceinline__ float qdec_lo(unsigned w, float cell, float org)
{
    return fmaf(__uint_as_float(__byte_perm(w, 0x4B000000u, 0x7410)) - 8388608.0f, cell, org);
}
__device__ __forceinline__ float qdec_hi(unsigned w, float cell, float org)
{
    return fmaf(__uint_as_float(__byte_perm(w, 0x4B000000u, 0x7432)) - 8388608.0f, cell, org);
}

// Slab distances of the 32-byte nodes straight from the grid coordinates: with, per ray and axis, ig = cell / d and
// cg = (org - o) / d, the distance to the plane with grid coordinate q is fma(q, ig, cg) -- a byte permute, an add (q as a
// float) and the fma: three instructions per plane instead of five (decode to world, subtract the origin, multiply).
// The product is exact inside the fma; ig and cg carry half an ulp each, i.e. the plane moves by less than 1e-2 of a cell
// -- the boxes are rounded outwards by a whole one.  (Folding the 2^23 of the conversion trick into cg as well -- two
// instructions per plane -- costs a whole cell of precision for every 1 / |d|, weakens the pruning of rays that run
// nearly parallel to an axis and measured 5 % SLOWER.)
// A direction component of exactly 0 must NOT become 1 / d = inf here: q * inf - inf = NaN, which fminf / fmaxf drop -- "no
// constraint" on that axis -- and a ray with TWO zero components then passes every box of its column and walks the whole
// tree.  Such rays exist: the cosine-hemisphere sample with u1 rounding to 1.0 (64 of the 2^31 generator states) leaves an
// axis-aligned surface exactly along its normal -- about one ray per launch at config 5's size, and its 3.5 ms walk was the
// tail that made this form 30-45 % slower on bounce 1 (ncu: one SM active 17.9 M cycles, the others 11.1 M,
// profiles/r2_ab_variants.txt).  1 / d is therefore clamped to +-1e20: the ray then drifts by 1e-20 t along that axis, and the
// rounding of the two terms (now about 1e21 each) moves a plane by less than 1e-5 -- far inside the outward cell of the boxes.
struct QRay {
    V3 ig, cg;
};
__device__ __forceinline__ QRay qray(const V3& o, const V3& inv_in, const V3& qorg, const V3& qcell)
{
    const V3 inv = v3(fminf(fmaxf(inv_in.x, -1e20f), 1e20f), fminf(fmaxf(inv_in.y, -1e20f), 1e20f), fminf(fmaxf(inv_in.z, -1e20f), 1e20f));
    QRay r;
    r.ig = v3(qcell.x * inv.x, qcell.y * inv.y, qcell.z * inv.z);
    r.cg = v3((qorg.x - o.x) * inv.x, (qorg.y - o.y) * inv.y, (qorg.z - o.z) * inv.z);
    return r;
}
__device__ __forceinline__ float qlo(unsigned w) { return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7410)) - 8388608.0f; }
__device__ __forceinline__ float qhi(unsigned w) { return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7432)) - 8388608.0f; }
// entry distance into the box whose planes lie at the distances (ax, bx) x (ay, by) x (az, bz), or negative on a miss
__device__ __forceinline__ float qbox_entry(float ax, float bx, float ay, float by, float az, float bz, float t_lim)
{
    const float tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.0f));
    const float tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), t_lim));
    return tn <= tf * 1.00001f + 1e-5f ? tn : -1.0f;
}

// conservative ray / box test of the geometry BVH: entry distance, or a negative value on a miss.
// Not parity-critical (it only prunes), so plain fast arithmetic; fminf / fmaxf drop the NaNs of
// 0 * inf, which makes a degenerate slab "no constraint".
__device__ __forceinline__ float gbox_entry(const float4& lo, const float4& hi, const V3& o, const V3& inv, float t_lim)
{
    const float ax = (lo.x - o.x) * inv.x, bx = (hi.x - o.x) * inv.x;
    const float ay = (lo.y - o.y) * inv.y, by = (hi.y - o.y) * inv.y;
    const float az = (lo.z - o.z) * inv.z, bz = (hi.z - o.z) * inv.z;
    const float tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.0f));
    const float tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), t_lim));
    return tn <= tf * 1.00001f + 1e-5f ? tn : -1.0f;
}

// closest hit among the sphere / cube primitives through the geometry BVH.
// Two alternating phases keep the warp converged although every thread walks its own way through
// the tree: (1) traversal with box tests only -- the primitives of a reached leaf are tested
// against their OWN boxes and the survivors appended to a per-thread candidate list, in roughly
// near-to-far order (keeping the list sorted costs more than it prunes: measured); (2) the exact tests of the listed candidates, where a thread first skips the
// candidates that start behind its current hit, so that every trip of the loop does real work in
// every lane that is still busy.  (Testing leaves as they are met costs a full exact test per
// trip as soon as ONE lane of the warp sits at a leaf.)
constexpr int PT_GCAND = 32;       // candidate list capacity; drained when fewer than 16 slots are free

__device__ __forceinline__ void closest_in_gbvh(const RenderParams& p, const float4* __restrict__ snodes_g, const V3& o,
                                                const V3& d, float& t_min, int& hit, Cand& best)
{
    const V3 inv = v3(__fdividef(1.0f, d.x), __fdividef(1.0f, d.y), __fdividef(1.0f, d.z));
    int stack[64];
    int cand_prim[PT_GCAND];
    float cand_t[PT_GCAND];
    int sp = 0, nc = 0, node = 0;
    bool more = true;
    while (more) {
        // ---- phase 1: traverse, collect candidates ----
        // the pruning distance trails the exact t_min by a margin far above the rounding of the box test
        const float t_lim = fmaf(t_min, 1.0001f, 1e-4f);
        while (true) {
            const float4* q = node < p.gb.n_smem_nodes ? snodes_g + 4 * node : p.gb.nodes + 4 * (size_t)node;
            const float4 q0 = q[0], q1 = q[1], q2 = q[2], q3 = q[3];
            const float tl = gbox_entry(q0, q1, o, inv, t_lim), tr = gbox_entry(q2, q3, o, inv, t_lim);
            int ref_near = __float_as_int(q0.w), ref_far = __float_as_int(q1.w);
            bool near_ok = tl >= 0.0f, far_ok = tr >= 0.0f;
            if (far_ok && (!near_ok || tr < tl)) {
                const int r = ref_near; ref_near = ref_far; ref_far = r;
                const bool b = near_ok; near_ok = far_ok; far_ok = b;
            }
            int next = -1;
#pragma unroll
            for (int side = 0; side < 2; ++side) {
                const int ref = side == 0 ? ref_near : ref_far;
                const bool ok = side == 0 ? near_ok : far_ok;
                if (!ok) continue;
                if (ref < 0) {
                    const int v = ~ref, first = v >> 3, count = (v & 7) + 1;
                    for (int k = 0; k < count; ++k) {
                        const float te = gbox_entry(p.gb.boxes[2 * (first + k)], p.gb.boxes[2 * (first + k) + 1], o, inv, t_lim);
                        if (te >= 0.0f) { cand_prim[nc] = first + k; cand_t[nc] = te; ++nc; }
                    }
                } else if (next < 0) {
                    next = ref;
                } else {
                    stack[sp++] = ref;
                }
            }
            if (next >= 0) node = next;
            else if (sp > 0) node = stack[--sp];
            else { more = false; break; }
            if (nc > PT_GCAND - 16) break;
        }
        // ---- phase 2: exact tests of the candidates ----
        int k = 0;
        while (true) {
            const float lim = fmaf(t_min, 1.0001f, 1e-4f);
            while (k < nc && cand_t[k] > lim) ++k;
            if (k >= nc) break;
            const PrimRec& g = p.gb.prims[cand_prim[k]];
            ++k;
            const Cand c = prim_test_t<true>(g, o, d);
            const int gi = g.pad0;
            // lexicographic minimum of (t, index): what the ascending loop with strict `<` finds
            if (c.t > 0.0f && (t_min > c.t || (t_min == c.t && gi < hit))) {
                t_min = c.t;
                hit = gi;
                best = c;
            }
        }
        nc = 0;
    }
}

// ---- small scenes: no tree, conservative per-primitive tests that also bound the distance ----------
// With a handful of primitives (Cornell: 7) walking a tree costs more than it saves, and what limits
// the kernel is how many EXACT tests a warp runs: the candidates of a ray used to be tested in index
// order, so a warp ran as many trips of the exact test as its unluckiest lane had overlapping boxes
// (ncu: 21 of 32 lanes active there, profiles/r4_*).  Here every primitive gets a cheap conservative
// test against its true shape (world box for an axis-aligned cube, object-space slabs for a rotated one,
// the quadratic for a sphere), all lanes on the same primitive at the same time, which yields a lower
// bound t_lo of the distance the reference would report (pt_gbvh.cu: margins d_hit / d_t).  The exact
// test then runs on the candidate with the smallest t_lo first -- nearly always the winner, after
// which the others are pruned by t_lo > t -- so most rays cost ONE exact test.  Selection is by the
// lexicographic minimum of (t, index), as everywhere.
__device__ __forceinline__ bool cull_test(const float4& q0, const float4& q1, const PrimRec& g, const V3& o, const V3& d,
                                          const V3& inv, float& t_lo)
{
    const int kind = __float_as_int(q0.w);
    float nx, ny, nz, fx, fy, fz, kx, ky, kz;
    if (kind == 0) {
        const float ax = (q0.x - o.x) * inv.x, bx = (q1.x - o.x) * inv.x;
        const float ay = (q0.y - o.y) * inv.y, by = (q1.y - o.y) * inv.y;
        const float az = (q0.z - o.z) * inv.z, bz = (q1.z - o.z) * inv.z;
        nx = fminf(ax, bx); fx = fmaxf(ax, bx);
        ny = fminf(ay, by); fy = fmaxf(ay, by);
        nz = fminf(az, bz); fz = fmaxf(az, bz);
        kx = q1.w * fabsf(inv.x); ky = q1.w * fabsf(inv.y); kz = q1.w * fabsf(inv.z);
    } else if (kind == 1 || kind == 2) {
        const float* m = g.inv;
        const V3 qo = v3(fmaf(o.x, m[0], fmaf(o.y, m[3], fmaf(o.z, m[6], m[9]))), fmaf(o.x, m[1], fmaf(o.y, m[4], fmaf(o.z, m[7], m[10]))),
                         fmaf(o.x, m[2], fmaf(o.y, m[5], fmaf(o.z, m[8], m[11]))));
        const V3 qd = v3(fmaf(d.x, m[0], fmaf(d.y, m[3], d.z * m[6])), fmaf(d.x, m[1], fmaf(d.y, m[4], d.z * m[7])),
                         fmaf(d.x, m[2], fmaf(d.y, m[5], d.z * m[8])));
        if (kind == 2) {                         // |qo + t qd|^2 = r^2; qd is not normalised, so t is the world distance
            const float a = fmaf(qd.x, qd.x, fmaf(qd.y, qd.y, qd.z * qd.z));
            const float b = fmaf(qo.x, qd.x, fmaf(qo.y, qd.y, qo.z * qd.z));
            const float c = fmaf(qo.x, qo.x, fmaf(qo.y, qo.y, qo.z * qo.z));
            // The reference's radicand dot(o,d)^2 - (dot(o,o) - r^2) (intersections.cu:74-79) cancels numbers of
            // size c = |qo|^2 and inherits the rounding q0.x of qo: the sphere it sees may be this much larger.
            // The margin grows with the distance of the origin, so a ray LEAVING the sphere (origin 1e-4 outside
            // it) keeps a tight one and is rejected here instead of costing an exact test.
            const float rh2 = 0.25f + fmaf(1.9073486e-6f, c + 0.25f, q0.x * fmaf(4.0f, sqrtf(c), 4.0f * q0.x));
            const float disc = fmaf(b, b, -a * (c - rh2));
            if (!(disc >= 0.0f)) return false;
            const float sq = sqrtf(disc);
            if (sq - b < -9.5367432e-7f * (sq + fabsf(b))) return false;          // the far root lies behind the origin
            const float ia = __fdividef(1.0f, a);
            const float rt = sqrtf(rh2) + q0.y;                                    // + d_t: its near root bounds t from below
            const float disc_t = fmaf(b, b, -a * (c - rt * rt * 1.000001f));
            t_lo = (-b - sqrtf(fmaxf(disc_t, disc))) * ia;
            return true;
        }
        const V3 iq = v3(__fdividef(1.0f, qd.x), __fdividef(1.0f, qd.y), __fdividef(1.0f, qd.z));
        const float ax = (-q0.x - qo.x) * iq.x, bx = (q0.x - qo.x) * iq.x;
        const float ay = (-q0.y - qo.y) * iq.y, by = (q0.y - qo.y) * iq.y;
        const float az = (-q0.z - qo.z) * iq.z, bz = (q0.z - qo.z) * iq.z;
        nx = fminf(ax, bx); fx = fmaxf(ax, bx);
        ny = fminf(ay, by); fy = fmaxf(ay, by);
        nz = fminf(az, bz); fz = fmaxf(az, bz);
        kx = q1.x * fabsf(iq.x); ky = q1.y * fabsf(iq.y); kz = q1.z * fabsf(iq.z);
    } else {
        t_lo = -3.402823466e+38f;
        return true;
    }
    // fminf / fmaxf drop the NaNs of 0 * inf: a degenerate slab is "no constraint"
    const float tn = fmaxf(fmaxf(nx, ny), nz), tf = fminf(fminf(fx, fy), fz);
    if (!(tn <= tf && tf >= 0.0f)) return false;
    t_lo = fmaxf(fmaxf(nx - kx, ny - ky), nz - kz);      // entry into the box grown by d_t
    return true;
}

#ifndef PT_CULL_FAST
#define PT_CULL_FAST 1
#endif
// sp: the n sphere / cube records (pad0 = index in the scene's primitive array), cr: their cull records
__device__ __forceinline__ void closest_small(const PrimRec* __restrict__ sp, const float4* __restrict__ cr, int n, const V3& o,
                                              const V3& d, float& t_min, int& hit, int& hit_k, Cand& best)
{
    const V3 inv = v3(__fdividef(1.0f, d.x), __fdividef(1.0f, d.y), __fdividef(1.0f, d.z));
    const float INF = 3.402823466e+38f;
    float t1 = INF, t2 = INF;                    // the two smallest t_lo
    int k1 = -1, k2 = -1;
    unsigned cand = 0u;
    // (a runtime loop on purpose: unrolling the first 4 / 8 steps lost 2 % / 7 % -- spills and instruction-cache misses,
    //  profiles/r2_ab_variants.txt)
#if PT_CULL_FAST
    // The record's address is a register that advances by 32 bytes, the record is read once (kind is its fourth word), and the
    // two smallest lower bounds are kept with selects on ALL lanes -- a primitive that is no candidate enters with +inf --
    // instead of a divergent branch that a handful of lanes took: 42 instead of 60 issue slots per primitive.
    unsigned cra = smem_u32(cr);
    for (int k = 0; k < n; ++k, cra += 32u) {
        const float4 q0 = lds_f4(cra), q1 = lds_f4(cra + 16u);
        float t_lo;
        const bool ok = cull_test(q0, q1, sp[k], o, d, inv, t_lo);
        t_lo = ok ? t_lo : INF;
        cand |= (ok ? 1u : 0u) << k;
        const bool lt1 = t_lo < t1, lt2 = t_lo < t2;
        k2 = lt1 ? k1 : (lt2 ? k : k2);
        t2 = lt1 ? t1 : (lt2 ? t_lo : t2);
        k1 = lt1 ? k : k1;
        t1 = lt1 ? t_lo : t1;
    }
#else
    for (int k = 0; k < n; ++k) {
        float t_lo;
        if (cull_test(cr[2 * k], cr[2 * k + 1], sp[k], o, d, inv, t_lo)) {
            cand |= 1u << k;
            if (t_lo < t1) { t2 = t1; k2 = k1; t1 = t_lo; k1 = k; }
            else if (t_lo < t2) { t2 = t_lo; k2 = k; }
        }
    }
#endif
    if (k1 < 0) return;
    cand &= ~(1u << k1);
    if (k2 >= 0) cand &= ~(1u << k2);
    int next = k1;
    bool first = true;
    while (next >= 0) {                          // one call site of the exact test: the kernel is large enough as it is
        const PrimRec& g = sp[next];
        const Cand c = prim_test(g, o, d);
        const int gi = g.pad0;
        if (c.t > 0.0f && (t_min > c.t || (t_min == c.t && gi < hit))) {
            t_min = c.t;
            hit = gi;
            hit_k = next;
            best = c;
        }
        if (first) {
            first = false;
            // no margin needed: t_lo is a true lower bound of what the reference would report (pt_gbvh.cu)
            if (k2 >= 0 && !(t2 > t_min)) { next = k2; continue; }
        }
        if (cand) {                              // three or more candidates: rare, no pruning
            next = __ffs(cand) - 1;
            cand &= cand - 1u;
        } else {
            next = -1;
        }
    }
}

// the MESH primitives: the candidate of each is the BVH hit if the hit triangle belongs to it
__device__ __forceinline__ void closest_among_mesh_prims(const RenderParams& p, const MeshHit& mh, float& t_min, int& hit,
                                                         int& owner)
{
    for (int k = 0; k < p.gb.n_mesh_prims; ++k) {
        const int gi = p.gb.mesh_prims[k];
        const PrimRec& g = p.prims[gi];
        if (mh.any && mh.id >= g.pad0 && mh.id < g.pad1) {
            if (owner < 0) owner = gi;
            if (mh.t > 0.0f && (t_min > mh.t || (t_min == mh.t && gi < hit))) {
                t_min = mh.t;
                hit = gi;
            }
        }
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(PT_THREADS) k_raygen(const __grid_constant__ RenderParams p)
{
    const int slot = blockIdx.y;
    const int idx = blockIdx.x * PT_THREADS + threadIdx.x;
    if (idx == 0) p.counts[slot] = p.P;             // live paths entering bounce 0
    if (idx == 0 && slot == 0)
        for (int b = 0; b < 2 * p.depth; ++b) p.tickets[b] = 0; // work-item counters of this batch's bounce launches ([depth..2 depth): closest hit)
    if (idx >= p.P) return;
    const int y = idx / p.cam.W, x = idx - y * p.cam.W;
    const int iter = p.iter0 + slot * p.stride;
    const PathState s = make_camera_path(p.cam, p.opt, iter, p.depth, x, y);
    const size_t at = (size_t)slot * p.P + idx;
    p.in.a[at] = make_float4(s.o.x, s.o.y, s.o.z, __int_as_float(s.pixel));
    p.in.b[at] = make_float4(s.d.x, s.d.y, s.d.z, __int_as_float(s.remaining));
    p.in.c[at] = make_float4(s.c.x, s.c.y, s.c.z, __int_as_float(s.bounces));
}

// the bookkeeping of k_raygen without the rays (bounce 0 of k_bounce generates them itself: RenderParams::raygen)
__global__ void __launch_bounds__(64) k_batch_init(const __grid_constant__ RenderParams p)
{
    for (int sl = threadIdx.x; sl < p.batch; sl += 64) p.counts[sl] = p.P;
    for (int b = threadIdx.x; b < 2 * p.depth; b += 64) p.tickets[b] = 0;
}

// ------------------------------------------------------------------------------------------------
// Pieces shared by the fused bounce kernel and the split (material-sorted) pipeline.

struct HitRec {       // ShadeableIntersection (sceneStructs.h:92-97)
    float t;          // -1 = miss
    V3 n;
    int mat;          // 0 on a miss: the reference memsets the record (pathtrace.cu:466)
    float u, v;       // texture coordinate (mesh hits; 0 otherwise)
};

__device__ __forceinline__ MeshHit no_mesh_hit()
{
    MeshHit h;
    h.any = false; h.t = 3.402823466e+38f; h.tri = -1; h.id = -1; h.bx = 0.f; h.by = 0.f;
    return h;
}

__device__ __forceinline__ void bounce_barriers_init(unsigned long long* bar)
{
    if (threadIdx.x == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
}

// producer (thread 0): stage the materials (optional) and the first one or two primitive chunks
// (with the geometry BVH: the top of that tree instead of primitive chunks)
__device__ __forceinline__ void stage_begin(const RenderParams& p, unsigned long long* bar, MatRec* smats, int n_smem_mats,
                                            BvhRec* snodes, PrimRec* sprims, bool want_prims, float4* scull = nullptr,
                                            int mesh_nodes = -1)
{
    if (threadIdx.x != 0) return;
    const int n_mesh_nodes = mesh_nodes >= 0 ? mesh_nodes : p.mesh.n_smem_nodes;     // nodes of the mesh BVH to stage
    if (p.gb.n_nodes > 0) {
        const unsigned mat_bytes = (unsigned)(n_smem_mats * sizeof(MatRec));
        const unsigned node_bytes = want_prims ? (unsigned)(n_mesh_nodes * sizeof(BvhRec)) : 0u;
        if (p.gb.n_small > 0) {                  // the records themselves and their cull records, no tree
            const unsigned r_bytes = want_prims ? (unsigned)(p.gb.n_small * sizeof(PrimRec)) : 0u;
            const unsigned c_bytes = want_prims ? (unsigned)(p.gb.n_small * 2 * sizeof(float4)) : 0u;
            mbar_expect_tx(&bar[0], mat_bytes + node_bytes + r_bytes + c_bytes);
            if (mat_bytes) bulk_g2s(smats, p.mats, mat_bytes, &bar[0]);
            if (node_bytes) bulk_g2s(snodes, p.mesh.nodes, node_bytes, &bar[0]);
            if (r_bytes) bulk_g2s(sprims, p.gb.prims, r_bytes, &bar[0]);
            if (c_bytes) bulk_g2s(scull ? scull : reinterpret_cast<float4*>(sprims + PT_SMALL_MAX), p.gb.crecs, c_bytes, &bar[0]);
            return;
        }
        const unsigned g_bytes = want_prims ? (unsigned)(p.gb.n_smem_nodes * 4 * sizeof(float4)) : 0u;
        mbar_expect_tx(&bar[0], mat_bytes + node_bytes + g_bytes);
        if (mat_bytes) bulk_g2s(smats, p.mats, mat_bytes, &bar[0]);
        if (node_bytes) bulk_g2s(snodes, p.mesh.nodes, node_bytes, &bar[0]);
        if (g_bytes) bulk_g2s(sprims, p.gb.nodes, g_bytes, &bar[0]);
        return;
    }
    const int n_chunks = (p.n_prims + PT_CHUNK - 1) / PT_CHUNK;
    const int chunk_cap = n_chunks > 1 ? PT_CHUNK : p.n_prims;
    const unsigned mat_bytes = (unsigned)(n_smem_mats * sizeof(MatRec));
    const unsigned node_bytes = want_prims ? (unsigned)(n_mesh_nodes * sizeof(BvhRec)) : 0u;
    const int n0 = want_prims ? min(chunk_cap, p.n_prims) : 0;
    mbar_expect_tx(&bar[0], mat_bytes + node_bytes + (unsigned)(n0 * sizeof(PrimRec)));
    if (mat_bytes) bulk_g2s(smats, p.mats, mat_bytes, &bar[0]);
    if (node_bytes) bulk_g2s(snodes, p.mesh.nodes, node_bytes, &bar[0]);
    if (n0) bulk_g2s(sprims, p.prims, (unsigned)(n0 * sizeof(PrimRec)), &bar[0]);
    if (want_prims && n_chunks > 1) {
        const int n1 = min(PT_CHUNK, p.n_prims - PT_CHUNK);
        mbar_expect_tx(&bar[1], (unsigned)(n1 * sizeof(PrimRec)));
        bulk_g2s(sprims + PT_CHUNK, p.prims + PT_CHUNK, (unsigned)(n1 * sizeof(PrimRec)), &bar[1]);
    }
}

// computeIntersections (pathtrace.cu:230-298) over the TMA-staged primitive chunks.  Every thread of
// the CTA must call it (it contains the chunk barriers); inactive threads only take part in those.
// GB: the sphere / cube primitives are searched through the geometry BVH (whose top levels lie
// where the primitive chunks would) instead of the linear scan.
// GB: 0 = linear scan over the staged primitive chunks, 1 = geometry BVH, 2 = small-scene search
// MESH = false: the caller knows that the scene has no triangle meshes (the persistent kernel's small-scene mode: launch_bounce
// sends a small scene WITH meshes to the tile kernel) -- the mesh walk, its visit stack and leaf lists and the triangle
// attributes are then not compiled into the kernel at all.
template <int GB, bool MESH = true>
__device__ __forceinline__ HitRec closest_hit_staged(const RenderParams& p, unsigned long long* bar, const BvhRec* snodes,
                                                     const PrimRec* sprims, bool active, const V3& o, const V3& d)
{
    const int n_chunks = GB ? 0 : (p.n_prims + PT_CHUNK - 1) / PT_CHUNK;
    float t_min = 3.402823466e+38f;
    int hit = -1;
    Cand best;
    best.t = -1.f; best.outside = true; best.aux = v3(0.f, 0.f, 0.f);
    MeshHit mh = no_mesh_hit(), last = no_mesh_hit();
    int owner = -1;
    if (GB || (MESH && p.mesh.n_nodes > 0)) mbar_wait(&bar[0], 0u);     // the staged top(s) of the tree(s)
    // A ray whose direction is NaN (glm::refract past the critical angle, interactions.cu:68) hits
    // nothing in the reference: every t it computes is NaN and `t > 0` fails.  In a tree such a ray
    // passes every box test (NaN compares false) and would visit all of it: skip the trees, same result.
    const bool walk = active && d.x == d.x && d.y == d.y && d.z == d.z;
    int hit_k = -1;
    if (GB) {
        // spheres / cubes first: their nearest hit caps the mesh walk (when nothing behind it can show)
        if (GB == 2) {
            if (walk) closest_small(sprims, reinterpret_cast<const float4*>(sprims + PT_SMALL_MAX), p.gb.n_small, o, d, t_min, hit, hit_k, best);
        } else if (walk) closest_in_gbvh(p, reinterpret_cast<const float4*>(sprims), o, d, t_min, hit, best);
        if (MESH && p.mesh.n_nodes > 0 && walk) {
            const float cap = p.gb.cap_mesh_walk ? prune_limit(t_min) : 3.402823466e+38f;
            mh = bvh_traverse(p.mesh, snodes, o, d, cap);       // one traversal serves every MESH primitive
            closest_among_mesh_prims(p, mh, t_min, hit, owner);
        }
    } else if (MESH && p.mesh.n_nodes > 0 && walk && !p.mesh.brute) {
        mh = bvh_traverse(p.mesh, snodes, o, d);
    }
    for (int ch = 0; ch < n_chunks; ++ch) {
        mbar_wait(&bar[ch & 1], (unsigned)((ch >> 1) & 1));
        const int n = min(PT_CHUNK, p.n_prims - ch * PT_CHUNK);
        const PrimRec* sp = sprims + (n_chunks > 1 ? (ch & 1) * PT_CHUNK : 0);
        if (active) closest_in_chunk(sp, n, ch * PT_CHUNK, o, d, p.mesh, mh, last, t_min, hit, best, owner);
        if (ch + 2 < n_chunks) {                     // refill this buffer with chunk ch + 2
            __syncthreads();
            if (threadIdx.x == 0) {
                const int n2 = min(PT_CHUNK, p.n_prims - (ch + 2) * PT_CHUNK);
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_expect_tx(&bar[ch & 1], (unsigned)(n2 * sizeof(PrimRec)));
                bulk_g2s(const_cast<PrimRec*>(sprims) + (ch & 1) * PT_CHUNK, p.prims + (size_t)(ch + 2) * PT_CHUNK,
                         (unsigned)(n2 * sizeof(PrimRec)), &bar[ch & 1]);
            }
        }
    }
    HitRec h;
    h.t = -1.0f;
    h.n = v3(0.f, 0.f, 0.f);
    h.mat = 0;
    h.u = h.v = 0.0f;
    if (hit >= 0) {
        // (small-scene search: a MESH winner is not among the staged records -> hit_k stays behind)
        const bool staged = GB == 2 ? (hit_k >= 0 && sprims[hit_k].pad0 == hit) : (GB == 0 && n_chunks <= 1);
        const PrimTail g = load_prim_tail(staged, smem_u32(sprims + (GB == 2 ? hit_k : hit)), p.prims + hit);
        if (MESH) hit_attributes(p.mesh, g, hit, best, mh, owner, h.n, h.u, h.v);
        else h.n = prim_normal(g.itr, g.type, best);
        h.t = t_min;
        h.mat = g.material;
    }
    return h;
}

__device__ __forceinline__ PathState load_path(const Pool& pool, size_t at)
{
    const float4 a = pool.a[at], b = pool.b[at], c = pool.c[at];
    PathState s;
    s.o = v3(a.x, a.y, a.z);  s.pixel = __float_as_int(a.w);
    s.d = v3(b.x, b.y, b.z);  s.remaining = __float_as_int(b.w);
    s.c = v3(c.x, c.y, c.z);  s.bounces = __float_as_int(c.w);
    return s;
}

// shadeMaterials + scatterRay for the path at compacted slot `idx`, finalGather for a path that
// ends here, and the stable single-pass compaction of the survivors into the other pool.
// Every thread of the CTA must call it.
#ifndef PT_EXPLICIT_MAT
#define PT_EXPLICIT_MAT 1
#endif
__device__ __forceinline__ bool shade_path(const RenderParams& p, const MatRec* smats, int n_smem_mats, PathState& s,
                                           const HitRec& h, bool active, int slot, int idx)
{
    bool live = false;
    if (active) {
#if PT_EXPLICIT_MAT
        const MatRec m = load_mat(n_smem_mats != 0, smem_u32(smats + h.mat), p.mats + h.mat);
#else
        const MatRec& m = n_smem_mats ? smats[h.mat] : p.mats[h.mat];
#endif
        const int iter = p.iter0 + slot * p.stride;
        const V3 base = (m.texture_id != -1 && h.t > 0.0f) ? tex_color(p.mesh, m.texture_id, h.u, h.v)
                                                           : v3(m.color[0], m.color[1], m.color[2]);
        live = shade_exact(s, h.t, h.n, m, base, iter, idx, p.opt.rr);
        if (!live && (s.c.x != 0.0f || s.c.y != 0.0f || s.c.z != 0.0f)) {
            // finalGather (pathtrace.cu:347-356) for this path, at termination: one add per pixel
            // per iteration, issued in iteration order when batch == 1
            float* px = p.image + 3 * (size_t)s.pixel;
            atomicAdd(px + 0, s.c.x);
            atomicAdd(px + 1, s.c.y);
            atomicAdd(px + 2, s.c.z);
        }
    }
    return live;
}

__device__ __forceinline__ void shade_and_compact(const RenderParams& p, BounceShared& sh, const MatRec* smats,
                                                  int n_smem_mats, PathState& s, const HitRec& h, bool active,
                                                  int slot, int tile, int idx, int n_live, int* count_out)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool live = shade_path(p, smats, n_smem_mats, s, h, active, slot, idx);
    // ---- stable compaction of the survivors (mark_valid + scan + keep, pathtrace.cu:359-407) ---
    const unsigned bal = __ballot_sync(FULL_MASK, live);
    const int wrank = __popc(bal & ((1u << lane) - 1u));
    if (lane == 0) sh.warp_cnt[warp] = __popc(bal);
    __syncthreads();
    int warp_prefix = 0, aggregate = 0;
#pragma unroll
    for (int w = 0; w < PT_THREADS / 32; ++w) {
        const int c = sh.warp_cnt[w];
        if (w < warp) warp_prefix += c;
        aggregate += c;
    }
    if (warp == 0) {
        unsigned long long* desc = p.desc + (size_t)slot * p.tiles * PT_DESC_STRIDE;
        const int e = b200::lb_exclusive_prefix_cfg<1, b200::LB_DEFAULT_BACKOFF_NS, PT_DESC_STRIDE>(
            desc, (unsigned)tile, p.epoch, aggregate, 0, lane);
        if (lane == 0) {
            sh.tile_exclusive = e;
            if (tile == (n_live - 1) / PT_THREADS) *count_out = e + aggregate;
        }
    }
    __syncthreads();
    if (live) {
        const size_t at = (size_t)slot * p.P + (size_t)(sh.tile_exclusive + warp_prefix + wrank);
        p.out.a[at] = make_float4(s.o.x, s.o.y, s.o.z, __int_as_float(s.pixel));
        p.out.b[at] = make_float4(s.d.x, s.d.y, s.d.z, __int_as_float(s.remaining));
        p.out.c[at] = make_float4(s.c.x, s.c.y, s.c.z, __int_as_float(s.bounces));
    }
}

// ------------------------------------------------------------------------------------------------
// k_bounce: the fused bounce.  CACHE (first-bounce cache, bounce 0 only, valid when the camera rays
// do not change between iterations, i.e. SSAA and DoF off): 0 = off, 1 = trace and fill the cache
// (slot 0 writes it), 2 = replace the closest-hit search by one 20-byte cache read per path.
template <int CACHE, int GB>
__global__ void __launch_bounds__(PT_THREADS) k_bounce_tile(const __grid_constant__ RenderParams p)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    BounceShared& sh = *reinterpret_cast<BounceShared*>(smem_raw);
    const int n_smem_mats = p.n_mats <= PT_SMEM_MATS ? p.n_mats : 0;
    MatRec* smats = reinterpret_cast<MatRec*>(smem_raw + sizeof(BounceShared));
    BvhRec* snodes = reinterpret_cast<BvhRec*>(smats + n_smem_mats);
    __shared__ SmallStageT<GB == 2> small_st;
    PrimRec* sprims = GB == 2 ? small_st.prims : reinterpret_cast<PrimRec*>(snodes + (CACHE != 2 ? p.mesh.n_smem_nodes : 0));

    const int slot = blockIdx.y;
    const int tile = blockIdx.x;
    const int n_live = p.counts[p.bounce * p.batch + slot];
    int* count_out = p.counts + (p.bounce + 1) * p.batch + slot;
    if (tile * PT_THREADS >= n_live) {              // whole tile beyond the live range
        if (tile == 0 && threadIdx.x == 0) *count_out = 0;
        return;
    }
    if (tile == 0 && threadIdx.x == 0) atomicAdd(p.segments, (unsigned long long)n_live);
    bounce_barriers_init(sh.bar);
    stage_begin(p, sh.bar, smats, n_smem_mats, snodes, sprims, CACHE != 2);

    // this thread's path (overlaps the bulk copies)
    const int idx = tile * PT_THREADS + threadIdx.x;
    const bool active = idx < n_live;
    PathState s = load_path(p.in, (size_t)slot * p.P + (active ? idx : 0));

    HitRec h;
    if (CACHE == 2) {
        const float4 cn = p.cache_n[active ? idx : 0];
        const float2 cuv = p.cache_uv[active ? idx : 0];
        h.t = cn.x; h.n = v3(cn.y, cn.z, cn.w);
        h.mat = p.cache_k[active ? idx : 0];
        h.u = cuv.x; h.v = cuv.y;
        mbar_wait(&sh.bar[0], 0u);                  // the materials
    } else {
        h = closest_hit_staged<GB>(p, sh.bar, snodes, sprims, active, s.o, s.d);
        if (CACHE == 1 && slot == 0 && active) {
            p.cache_n[idx] = make_float4(h.t, h.n.x, h.n.y, h.n.z);
            p.cache_k[idx] = h.mat;
            p.cache_uv[idx] = make_float2(h.u, h.v);
        }
    }
    shade_and_compact(p, sh, smats, n_smem_mats, s, h, active, slot, tile, idx, n_live, count_out);
}

// ------------------------------------------------------------------------------------------------
// k_bounce: the fused bounce as a PERSISTENT kernel with a warp-specialised look-back.
//
// In k_bounce_tile every tile's 256 threads sit at a barrier while warp 0 looks back over the
// predecessors' descriptors -- and the predecessors are tiles that started at about the same time
// and take about as long, so the wait is a sizeable part of a tile's life (ncu, Cornell 800x800:
// 23-29 % of all warp samples were that one barrier, profiles/r4_k_bounce_*).  Here
//   * a CTA is 7 compute warps + 1 scan warp and keeps taking tiles (224 paths) from an atomic
//     ticket counter -- so a tile's predecessors are
//     always held by running CTAs that work their tickets off in order: no deadlock whatever the
//     residency -- with the scene staged into its shared memory by TMA once, not once per tile;
//   * the compute warps of tile k rank their survivors (ballot), stash them in shared memory (each
//     warp in its own 32-slot region), hand {slot, tile, aggregate} to the scan warp through a
//     two-entry mailbox, and go straight on to tile k+1;
//   * the scan warp publishes the aggregate, looks back, publishes the inclusive prefix (and the
//     live count of the next bounce after a slot's last tile) and posts the exclusive prefix;
//   * the compute warps pick that prefix up one tile later and write the stashed survivors of
//     tile k to the other pool, warp by warp (stable: e + ranks inside the tile).
// One CTA-wide barrier per tile is left (the seven warp counts), on named barrier 1 so that the
// scan warp is not part of it.
struct alignas(16) PersistShared {
    unsigned long long bar[2];            // TMA staging
    // Mailbox signalling between the compute warps and the scan warp: one single-arrival mbarrier per
    // parity and direction (waiting on an mbarrier suspends the warp in hardware; polling a flag in shared
    // memory cost 12 % of all issued instructions).  Request `it` uses req_bar[it & 1], phase it >> 1, and
    // cannot be posted before the answer to request it - 2 has been consumed, so a waiter never sees the
    // barrier more than one phase ahead.
    unsigned long long req_bar[2], resp_bar[2];
    int warp_cnt[2][PT_PWARPS + 1];       // survivors per compute warp, by tile parity
    int item[2][4];                       // {slot (-1: no more work), tile, -, -} of the tile with this parity
    int req[2][4];                        // mailbox to the scan warp: {slot, tile (-1: stop), aggregate, last tile of the slot}
    int resp[2];                          // exclusive prefix of the tile with this parity
    int total_items;
    int pad_;
    int n_live[PT_MAX_BATCH];             // live paths of each slot entering this bounce
    int item_base[PT_MAX_BATCH + 4];      // first work item of each slot (exclusive prefix of the live tile counts)
};
static_assert(sizeof(PersistShared) % 16 == 0, "the TMA destinations behind it must stay 16-byte aligned");

// Four CTAs per SM (64 registers; ptxas spills 12 words into the otherwise unused local frame) instead of the
// three that the natural 74 registers allow: +4.8 % path segments/s on the Cornell box (profiles/r5_ab_variants.txt).
// A cp.async prefetch of the next tile's paths, tried at the same time, changed nothing at 3 CTAs and lost 1 % at 4.
#ifndef PT_BOUNCE_MIN_CTAS
#define PT_BOUNCE_MIN_CTAS (1024 / PT_BTHREADS)
#endif
#ifndef PT_NO_MESH_SPECIAL
#define PT_NO_MESH_SPECIAL 1      // the persistent kernel's small-scene mode is compiled without the mesh walk (see closest_hit_staged)
#endif
#ifndef PT_PERM_PREFETCH
#define PT_PERM_PREFETCH 1
#endif
#ifndef PT_LB_K
#define PT_LB_K 1
#endif
constexpr size_t PT_STASH_BYTES = (size_t)3 * PT_PTILE * sizeof(float4);


// work item w -> (slot, tile); one thread
__device__ __forceinline__ void put_item(PersistShared& sh, int w, int parity)
{
    int slot = -1, tile = 0;
    if (w < sh.total_items) {
        slot = 0;
        while (w >= sh.item_base[slot + 1]) ++slot;
        tile = w - sh.item_base[slot];
    }
    sh.item[parity][0] = slot;
    sh.item[parity][1] = tile;
}

template <int CACHE, int GB>
__global__ void __launch_bounds__(PT_BTHREADS, PT_BOUNCE_MIN_CTAS) k_bounce(const __grid_constant__ RenderParams p)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    PersistShared& sh = *reinterpret_cast<PersistShared*>(smem_raw);
    const int n_smem_mats = p.n_mats <= PT_SMEM_MATS ? p.n_mats : 0;
    float4* stash = reinterpret_cast<float4*>(smem_raw + sizeof(PersistShared));      // [3][PT_PWARPS][32]
    MatRec* smats = reinterpret_cast<MatRec*>(smem_raw + sizeof(PersistShared) + PT_STASH_BYTES);
    BvhRec* snodes = reinterpret_cast<BvhRec*>(smats + n_smem_mats);
    __shared__ SmallStageT<GB == 2> small_st;
    PrimRec* sprims = GB == 2 ? small_st.prims : reinterpret_cast<PrimRec*>(snodes + (CACHE < 2 ? p.mesh.n_smem_nodes : 0));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // ---- prologue: the work-item table of this launch, the first ticket, the staging copies ----
    if (threadIdx.x < p.batch) sh.n_live[threadIdx.x] = p.counts[p.bounce * p.batch + threadIdx.x];
    if (threadIdx.x == 0) {
        mbar_init(&sh.req_bar[0], 1); mbar_init(&sh.req_bar[1], 1);
        mbar_init(&sh.resp_bar[0], 1); mbar_init(&sh.resp_bar[1], 1);
    }
    bounce_barriers_init(sh.bar);                    // fences the initialisations; contains a __syncthreads
    if (threadIdx.x == 0) {
        int base = 0;
        for (int sl = 0; sl < p.batch; ++sl) {
            sh.item_base[sl] = base;
            base += (sh.n_live[sl] + PT_PTILE - 1) / PT_PTILE;
        }
        sh.item_base[p.batch] = base;
        sh.total_items = base;
        put_item(sh, atomicAdd(p.tickets + p.bounce, 1), 0);
        // a CTA without work must not leave bulk copies in flight when it exits
        if (sh.item[0][0] >= 0) stage_begin(p, sh.bar, smats, n_smem_mats, snodes, sprims, CACHE < 2);
    }
    if (blockIdx.x == 0 && threadIdx.x < p.batch) {
        const int n = sh.n_live[threadIdx.x];
        if (n == 0) p.counts[(p.bounce + 1) * p.batch + threadIdx.x] = 0;       // nobody else writes a dead slot's count
        // (split pipeline: the closest-hit kernel has counted the segments, unless the first-bounce cache replaced it)
        else if (CACHE != 3 || p.cache_mode == 2) atomicAdd(p.segments, (unsigned long long)n);
    }
    __syncthreads();

    // ---- the scan warp ----
    if (warp == PT_PWARPS) {
        for (int it = 0;; ++it) {
            mbar_wait(&sh.req_bar[it & 1], (unsigned)((it >> 1) & 1));
            const int slot = sh.req[it & 1][0], tile = sh.req[it & 1][1];
            const int aggregate = sh.req[it & 1][2], last = sh.req[it & 1][3];
            if (tile < 0) break;
            unsigned long long* desc = p.desc + (size_t)slot * p.ptiles * PT_DESC_STRIDE;
            const int e = b200::lb_exclusive_prefix_cfg<PT_LB_K, b200::LB_DEFAULT_BACKOFF_NS, PT_DESC_STRIDE>(
                desc, (unsigned)tile, p.epoch, aggregate, 0, lane);
            __syncwarp();                             // every lane has read the request before it can be answered
            if (lane == 0) {
                if (last) p.counts[(p.bounce + 1) * p.batch + slot] = e + aggregate;
                sh.resp[it & 1] = e;
                mbar_arrive(&sh.resp_bar[it & 1]);
            }
        }
        return;
    }

    // ---- the compute warps ----
    float4* my_stash = stash + warp * 32;            // + k * PT_PTILE for word k of the path
    int prev_cnt = 0, prev_off = 0, prev_slot = 0;
    int perm_next = 0;                               // CACHE == 3: this thread's permutation entry of the NEXT tile
    int it = 0;
    for (;; ++it) {
        const int slot = sh.item[it & 1][0], tile = sh.item[it & 1][1];    // written before the previous tile's barrier
        if (slot < 0) break;                          // uniform over the CTA
        const int n_live = sh.n_live[slot];
        const int idx = tile * PT_PTILE + threadIdx.x;
        const bool active = idx < n_live;
        // CACHE == 3: slot idx of the (material-sorted) order is source slot perm[idx]; the RNG seed uses idx, exactly as
        // the reference's shading kernel indexes the sorted arrays (pathtrace.cu:479-497)
        int src = active ? idx : 0;
        // (the permutation entry of every tile but the CTA's first was fetched behind the previous tile's barrier: the gather
        //  below is then ONE dependent memory round trip per tile instead of two -- 45 % of this mode's stall samples were
        //  long-scoreboard waits on that chain, profiles/r2x_config4_k_bounce_shade_lines.txt)
        if (CACHE == 3 && active && p.perm_in)
            src = (PT_PERM_PREFETCH && it > 0) ? perm_next : p.perm_in[(p.perm_shared ? (size_t)0 : (size_t)slot * p.P) + idx];
        // Bounce 0 with the cache off: the camera path is computed here (generateRayFromCamera, pathtrace.cu:184-227, the same
        // function k_raygen calls) instead of being written by k_raygen and read back -- 96 B per path of HBM traffic and a
        // launch that was 5.3 % of a Cornell step (profiles/r2z_bench_launches.csv).
        PathState s;
        if (CACHE == 0 && p.raygen) {
            const int y = src / p.cam.W, x = src - y * p.cam.W;
            s = make_camera_path(p.cam, p.opt, p.iter0 + slot * p.stride, p.depth, x, y);
        } else {
            s = load_path(p.in, (size_t)slot * p.P + src);
        }

        HitRec h;
        if (CACHE == 3) {
            const size_t at = (size_t)slot * p.P + src;
            const float4 hn = p.cache_mode == 2 ? p.cache_n[src] : p.hit_n[at];
            const float2 huv = p.cache_mode == 2 ? p.cache_uv[src] : p.hit_uv[at];
            h.t = hn.x; h.n = v3(hn.y, hn.z, hn.w);
            h.mat = p.cache_mode == 2 ? p.cache_k[src] : (p.hit_k[at] & 0x00ffffff);     // (top byte: in-tile rank, k_isect)
            h.u = huv.x; h.v = huv.y;
            mbar_wait(&sh.bar[0], 0u);                // the materials
        } else if (CACHE == 2) {
            const float4 cn = p.cache_n[active ? idx : 0];
            const float2 cuv = p.cache_uv[active ? idx : 0];
            h.t = cn.x; h.n = v3(cn.y, cn.z, cn.w);
            h.mat = p.cache_k[active ? idx : 0];
            h.u = cuv.x; h.v = cuv.y;
            mbar_wait(&sh.bar[0], 0u);                // the materials
        } else {
            h = closest_hit_staged<GB, (GB != 2) || !PT_NO_MESH_SPECIAL>(p, sh.bar, snodes, sprims, active, s.o, s.d);
            if (CACHE == 1 && slot == 0 && active) {
                p.cache_n[idx] = make_float4(h.t, h.n.x, h.n.y, h.n.z);
                p.cache_k[idx] = h.mat;
                p.cache_uv[idx] = make_float2(h.u, h.v);
            }
        }
        // The CTA's next work item: the ticket is drawn here, between closest hit and shading, so that the
        // atomic's round trip hides behind the shading, and is turned into (slot, tile) after the barrier.
        // Not earlier: a ticket that sits unprocessed while higher ones are worked off elsewhere makes
        // every look-back wait for the straggler (drawn a whole tile ahead, half of all flushes found their
        // prefix not ready, 10 us each); not later: the whole CTA would wait for the round trip every tile.
        int next_w = 0;
        if (threadIdx.x == 0) next_w = atomicAdd(p.tickets + p.bounce, 1);
        const bool live = shade_path(p, smats, n_smem_mats, s, h, active, slot, idx);

        const unsigned bal = __ballot_sync(FULL_MASK, live);
        const int wrank = __popc(bal & ((1u << lane) - 1u));
        const int cnt = __popc(bal);
        if (lane == 0) sh.warp_cnt[it & 1][warp] = cnt;
        if (threadIdx.x == 0) put_item(sh, next_w, (it + 1) & 1);
        asm volatile("bar.sync 1, %0;" ::"n"(PT_PTILE) : "memory");
        if (PT_PERM_PREFETCH && CACHE == 3 && p.perm_in) {       // the next tile is known now: its permutation entry
            const int nslot = sh.item[(it + 1) & 1][0];
            if (nslot >= 0) {
                const int nidx = sh.item[(it + 1) & 1][1] * PT_PTILE + threadIdx.x;
                if (nidx < sh.n_live[nslot]) perm_next = p.perm_in[(p.perm_shared ? (size_t)0 : (size_t)nslot * p.P) + nidx];
            }
        }
        int warp_prefix = 0, aggregate = 0;
#pragma unroll
        for (int w = 0; w < PT_PWARPS; ++w) {
            const int c = sh.warp_cnt[it & 1][w];
            if (w < warp) warp_prefix += c;
            aggregate += c;
        }
        if (threadIdx.x == 0) {
            sh.req[it & 1][0] = slot;
            sh.req[it & 1][1] = tile;
            sh.req[it & 1][2] = aggregate;
            sh.req[it & 1][3] = tile == (n_live - 1) / PT_PTILE;
            mbar_arrive(&sh.req_bar[it & 1]);
        }
        if (it > 0) {                                 // the previous tile's survivors, now that its prefix is known
            mbar_wait(&sh.resp_bar[(it - 1) & 1], (unsigned)(((it - 1) >> 1) & 1));
            const int e = sh.resp[(it - 1) & 1];
            if (lane < prev_cnt) {
                const size_t at = (size_t)prev_slot * p.P + (size_t)(e + prev_off + lane);
                p.out.a[at] = my_stash[lane];
                p.out.b[at] = my_stash[PT_PTILE + lane];
                p.out.c[at] = my_stash[2 * PT_PTILE + lane];
            }
            __syncwarp();
        }
        if (live) {
            my_stash[wrank] = make_float4(s.o.x, s.o.y, s.o.z, __int_as_float(s.pixel));
            my_stash[PT_PTILE + wrank] = make_float4(s.d.x, s.d.y, s.d.z, __int_as_float(s.remaining));
            my_stash[2 * PT_PTILE + wrank] = make_float4(s.c.x, s.c.y, s.c.z, __int_as_float(s.bounces));
        }
        __syncwarp();
        prev_cnt = cnt;
        prev_off = warp_prefix;
        prev_slot = slot;
    }
    if (threadIdx.x == 0) {                           // tell the scan warp to leave
        sh.req[it & 1][1] = -1;
        mbar_arrive(&sh.req_bar[it & 1]);
    }
    if (it > 0) {
        mbar_wait(&sh.resp_bar[(it - 1) & 1], (unsigned)(((it - 1) >> 1) & 1));
        const int e = sh.resp[(it - 1) & 1];
        if (lane < prev_cnt) {
            const size_t at = (size_t)prev_slot * p.P + (size_t)(e + prev_off + lane);
            p.out.a[at] = my_stash[lane];
            p.out.b[at] = my_stash[PT_PTILE + lane];
            p.out.c[at] = my_stash[2 * PT_PTILE + lane];
        }
    }
}

struct SortShared {
    int cnt[PT_THREADS / 32][256];      // per-warp digit counts
};

// per-warp counts of each digit and this thread's stable rank among the equal digits of its warp
__device__ __forceinline__ int warp_digit_rank(SortShared& sh, bool active, int digit, int bins)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int b = lane; b < bins; b += 32) sh.cnt[warp][b] = 0;
    __syncwarp();
    const unsigned act = __ballot_sync(FULL_MASK, active);
    int rank = 0;
    if (active) {
        const unsigned peers = __match_any_sync(act, digit);
        rank = __popc(peers & ((1u << lane) - 1u));
        if (rank == 0) sh.cnt[warp][digit] = __popc(peers);
    }
    return rank;
}

// ------------------------------------------------------------------------------------------------
// The split pipeline used when paths are bucketed by material between intersection and shading
// (thrust::sort_by_key on materialId, pathtrace.cu:410-414,479-491): k_isect writes the hit
// records, k_sort_hist / (scan) / k_sort_scatter build the stable sorted order as a permutation
// (LSD radix, 8 bits per pass, one pass for up to 256 materials), k_shade gathers through it.
//
// With the material sort on and one radix pass (p.sort_bins > 0) k_isect also does the COUNTING half of the sort -- the
// material id is in a register here: it writes the tile's digit histogram column (tiles beyond a slot's live range write
// a zero column, as k_sort_hist does) and packs each path's STABLE RANK AMONG THE EQUAL KEYS OF ITS TILE into the top
// byte of the stored key, so that the scatter after the scan is a pure streaming pass (k_sort_scatter_packed:
// offset[key][tile] + rank; no warp match, no shared memory, no barrier).  Config 4: k_isect 16.3 -> 18.x ms instead of
// + 7.5 ms of k_sort_hist, k_sort_scatter 6.4 -> 3.6 ms (profiles/r2_ab_variants.txt).  A persistent form of this kernel
// (resident CTAs striding over the tiles, scene staged once per CTA) was measured 18 % SLOWER than one tile per CTA.
#ifndef PT_ISECT_TILES
#define PT_ISECT_TILES 16
#endif
#ifndef PT_ISECT_STRIDE
#define PT_ISECT_STRIDE 1
#endif
// A CTA takes PT_ISECT_TILES tiles of its slot, strided by the grid (the scene is staged once).  The grid covers every tile of every
// slot whatever the live counts are (they are device-side), and in the late bounces nearly all CTAs find their tile beyond the
// live range: with one tile per CTA config 4 launched 129 600 CTAs per bounce, ~100 us of CTA launches for a few microseconds
// of work -- a quarter of this kernel's time over a step (profiles/r2z_config4_launches.csv).  Strided, not consecutive:
// the live tiles of a slot (its first ones) then spread evenly over the slot's CTAs (config 4, spp/s: one tile per CTA 2870;
// 4 / 8 / 16 consecutive tiles 2946 / 2883 / 2720; 4 / 8 / 16 strided 3005 / 3044 / 3047).
template <int GB>
__host__ __device__ constexpr int isect_tiles_per_cta() { return GB == 0 ? 1 : PT_ISECT_TILES; }     // (GB == 0 re-arms its chunk barriers per tile)

template <int GB, bool MESH = true>
__global__ void __launch_bounds__(PT_THREADS, 4) k_isect(const __grid_constant__ RenderParams p)    // (64 registers: four CTAs per SM)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    BounceShared& sh = *reinterpret_cast<BounceShared*>(smem_raw);
    BvhRec* snodes = reinterpret_cast<BvhRec*>(smem_raw + sizeof(BounceShared));
    __shared__ SmallStageT<GB == 2> small_st;
    __shared__ int cnt[PT_THREADS / 32][256];
    PrimRec* sprims = GB == 2 ? small_st.prims : reinterpret_cast<PrimRec*>(snodes + p.mesh.n_smem_nodes);
    const int slot = blockIdx.y;
    const int n_live = p.counts[p.bounce * p.batch + slot];
    const int bins = p.sort_bins;
    constexpr int T = isect_tiles_per_cta<GB>();
    bool staged = false;
    for (int tt = 0; tt < T; ++tt) {
#if PT_ISECT_STRIDE
        const int tile = blockIdx.x + tt * gridDim.x;    // strided: the live tiles of a slot spread evenly over its CTAs
#else
        const int tile = blockIdx.x * T + tt;
#endif
        if (tile >= p.tiles) break;
        if (tile * PT_THREADS >= n_live) {               // (uniform over the CTA; the live tiles of a slot come first)
            int* hist = p.hist + (size_t)slot * bins * p.tiles;
            for (int bb = threadIdx.x; bb < bins; bb += PT_THREADS) hist[(size_t)bb * p.tiles + tile] = 0;
            continue;
        }
        if (tile == 0 && threadIdx.x == 0) atomicAdd(p.segments, (unsigned long long)n_live);
        if (!staged) {
            bounce_barriers_init(sh.bar);
            stage_begin(p, sh.bar, nullptr, 0, snodes, sprims, true);
            staged = true;
        }
        const int idx = tile * PT_THREADS + threadIdx.x;
        const bool active = idx < n_live;
        const size_t at = (size_t)slot * p.P + (active ? idx : 0);
        const float4 a = p.in.a[at], b = p.in.b[at];
        const HitRec h = closest_hit_staged<GB, MESH>(p, sh.bar, snodes, sprims, active, v3(a.x, a.y, a.z), v3(b.x, b.y, b.z));
        if (active) {
            p.hit_n[at] = make_float4(h.t, h.n.x, h.n.y, h.n.z);
            if (bins == 0) p.hit_k[at] = h.mat;
            p.hit_uv[at] = make_float2(h.u, h.v);
            if (p.cache_mode == 1 && slot == 0) {
                p.cache_n[idx] = make_float4(h.t, h.n.x, h.n.y, h.n.z);
                p.cache_k[idx] = h.mat;
                p.cache_uv[idx] = make_float2(h.u, h.v);
            }
        }
        if (bins > 0) {
            const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
            for (int bb = lane; bb < bins; bb += 32) cnt[warp][bb] = 0;
            __syncwarp();
            const unsigned act = __ballot_sync(FULL_MASK, active);
            int rank = 0;
            if (active) {
                const unsigned peers = __match_any_sync(act, h.mat & 255);
                rank = __popc(peers & ((1u << lane) - 1u));
                if (rank == 0) cnt[warp][h.mat & 255] = __popc(peers);
            }
            __syncthreads();
            if (active) {
                for (int ww = 0; ww < warp; ++ww) rank += cnt[ww][h.mat & 255];
                p.hit_k[at] = h.mat | (rank << 24);
            }
            int* hist = p.hist + (size_t)slot * bins * p.tiles;
            for (int bb = threadIdx.x; bb < bins; bb += PT_THREADS) {
                int t = 0;
#pragma unroll
                for (int ww = 0; ww < PT_THREADS / 32; ++ww) t += cnt[ww][bb];
                hist[(size_t)bb * p.tiles + tile] = t;
            }
            if (T > 1) __syncthreads();                  // the counts are read: the next tile may clear them
        }
    }
}

// Four paths per thread (one 16-byte load of their keys, four independent offset look-ups and stores): with one path per
// thread the pass ran at the latency of its three dependent memory operations -- 1 TB/s of keys at full occupancy (config 4:
// 3.6 ms of a 38 ms step).  A CTA takes four consecutive tiles per step, 64 threads each.
__global__ void __launch_bounds__(PT_THREADS) k_sort_scatter_packed(const __grid_constant__ RenderParams p, int bins)
{
    static_assert(PT_THREADS == 256, "four tiles of 256 paths per CTA step, 64 threads x 4 paths each");
    const int groups = (p.tiles + 3) / 4;
    const int items = p.batch * groups;
    const bool vec = (p.P & 3) == 0;                             // slot * P keeps the key array 16-byte aligned
    for (int w = blockIdx.x; w < items; w += gridDim.x) {
        const int slot = w / groups, g = w - slot * groups;
        const int n_live = p.counts[p.bounce * p.batch + slot];
        const int tile = 4 * g + (threadIdx.x >> 6);
        const int j0 = tile * PT_THREADS + (threadIdx.x & 63) * 4;
        if (j0 >= n_live) continue;                              // (n_live <= tiles * 256: also a tile past the last one)
        const int* offs = p.hist_scanned + (size_t)slot * bins * p.tiles;
        const int base0 = offs[0];
        const int* keys = p.hit_k + (size_t)slot * p.P + j0;
        int key[4];
        if (vec && j0 + 3 < n_live) {
            const int4 k4 = *reinterpret_cast<const int4*>(keys);
            key[0] = k4.x; key[1] = k4.y; key[2] = k4.z; key[3] = k4.w;
        } else {
#pragma unroll
            for (int q = 0; q < 4; ++q) key[q] = j0 + q < n_live ? keys[q] : 0;
        }
        int pos[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) pos[q] = offs[(size_t)(key[q] & 255) * p.tiles + tile] - base0 + (int)((unsigned)key[q] >> 24);
        int* out = p.perm_out + (size_t)slot * p.P;
#pragma unroll
        for (int q = 0; q < 4; ++q) if (j0 + q < n_live) out[pos[q]] = j0 + q;
    }
}

// ------------------------------------------------------------------------------------------------
// k_isect_dyn: closest hit through the geometry BVH with DYNAMIC RAY FETCH (mesh-free scenes).
//
// After the first bounce the rays of a warp have nothing in common, and how far each one walks the
// tree varies by orders of magnitude: in k_isect<1> a warp keeps running until its unluckiest ray is done
// (ncu, config 5, bounce 1: 6 of 32 threads active per instruction in the walk, profiles/r19_config5_*).
// The hit record of a ray goes to the ray's own index, so -- unlike shading, whose random numbers are
// seeded by the slot -- closest hit may take the rays in ANY order.  Here the kernel is persistent, every
// warp is its own worker, and a lane whose ray is finished gets a new one from a global ticket counter
// as soon as PT_DYN_REFILL lanes of the warp are idle; the lanes in between keep their state (stack,
// candidate list, current hit) in registers / local memory.  A round is: refill, up to PT_DYN_STEPS node
// visits per lane, the exact tests of the listed candidates once some lane needs them (list nearly full or
// walk finished; every lane with candidates joins in), write-out of finished rays.  The search itself --
// box tests, candidate pruning, lexicographic (t, index) winner -- is closest_in_gbvh's.
#ifndef PT_DYN_REFILL
#define PT_DYN_REFILL 8
#endif
#ifndef PT_DYN_STEPS
#define PT_DYN_STEPS 8
#endif
#ifndef PT_DYN_QFMA
#define PT_DYN_QFMA 1              // slab distances as fma(q, cell / d, (org - o) / d) straight from the grid coordinates
#endif
#ifndef PT_DYN_CULL
#define PT_DYN_CULL 0
#endif
#ifndef PT_NODE_SPLIT
#define PT_NODE_SPLIT 1
#endif
#ifndef PT_DYN_GNODES
#define PT_DYN_GNODES 0            // staged nodes of the quantised tree: none -- every lane loads its node from global memory through L1 (one code path;
                                   // 16 staged: 221.5 spp/s, none: 224.6; round 1: 512: 137, 64: 142, 16: 144)
#endif

struct DynShared {
    unsigned long long bar[2];
    int total;                        // live rays of all slots entering this bounce
    int pad_;
    int base[PT_MAX_BATCH + 2];       // exclusive prefix of the slots' live counts
};
#ifndef PT_DYN_POOL
#define PT_DYN_POOL 0                 // 1: warp-cooperative exact tests (k_isect_dyn) -- measured slower, see the A/B record
#endif
#ifndef PT_DYN_DRAIN_LANES
#define PT_DYN_DRAIN_LANES 1          // lanes that must need their candidates tested before the warp does it
#endif
struct DynWarpShared {                // per warp: the dealing pool and the way back of the results
    unsigned pool[PT_THREADS / 32][32];
    unsigned long long res[PT_THREADS / 32][32];
    float4 aux[PT_THREADS / 32][32];
};

__global__ void __launch_bounds__(PT_THREADS) k_isect_dyn(const __grid_constant__ RenderParams p)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    DynShared& sh = *reinterpret_cast<DynShared*>(smem_raw);
    const uint4* snodes_q = reinterpret_cast<const uint4*>(smem_raw + sizeof(DynShared));
    const int lane = threadIdx.x & 31;
    const int n_staged = p.gb.n_nodes < PT_DYN_GNODES ? p.gb.n_nodes : PT_DYN_GNODES;
    const unsigned snodes_s = smem_u32(snodes_q);
    const V3 qorg = v3(p.gb.qorg[0], p.gb.qorg[1], p.gb.qorg[2]), qcell = v3(p.gb.qcell[0], p.gb.qcell[1], p.gb.qcell[2]);
    bounce_barriers_init(sh.bar);
    if (threadIdx.x == 0) {
        int run = 0;
        for (int sl = 0; sl < p.batch; ++sl) {
            sh.base[sl] = run;
            run += p.counts[p.bounce * p.batch + sl];
        }
        sh.base[p.batch] = run;
        sh.total = run;
        if (blockIdx.x == 0) atomicAdd(p.segments, (unsigned long long)run);
        const unsigned bytes = (unsigned)(n_staged * 2 * sizeof(uint4));
        if (bytes) {
            mbar_expect_tx(&sh.bar[0], bytes);
            bulk_g2s(const_cast<uint4*>(snodes_q), p.gb.qnodes, bytes, &sh.bar[0]);
        } else {
            mbar_arrive(&sh.bar[0]);
        }
    }
    __syncthreads();
    mbar_wait(&sh.bar[0], 0u);
    const int total = sh.total;
    int* ticket = p.tickets + p.depth + p.bounce;       // (the shade launch of this bounce draws from tickets[bounce])

    // per-lane state of the ray in flight
    bool has = false, walking = false;
    int slot = 0, idx = 0;
    V3 o = v3(0.f, 0.f, 0.f), d = o, inv = o;
    QRay qr;
    qr.ig = o; qr.cg = o;
    float t_min = 3.402823466e+38f;
    int hit = -1;
    Cand best;
    best.t = -1.f; best.outside = true; best.aux = v3(0.f, 0.f, 0.f);
    static_assert(PT_GBVH_LEAF == 1, "the walk below takes a negative child reference for ONE primitive");
#if PT_DYN_POOL
    __shared__ DynWarpShared wsh;
#endif
    int stack[64];
    int cand_prim[PT_GCAND];
    float cand_t[PT_GCAND];
    int sp = 0, nc = 0, node = 0, top = 0;            // top: the newest stack entry (entries below it are in stack[0 .. sp-2])
    bool exhausted = false;                           // warp-uniform: the ticket counter has run past the last ray

    for (;;) {
        // ---- refill ----
        unsigned idle = __ballot_sync(FULL_MASK, !has);
        if (!exhausted && __popc(idle) >= PT_DYN_REFILL) {
            const int n = __popc(idle);
            int base = 0;
            if (lane == 0) base = atomicAdd(ticket, n);
            base = __shfl_sync(FULL_MASK, base, 0);
            if (!has) {
                const int w = base + __popc(idle & ((1u << lane) - 1u));
                if (w < total) {
                    slot = 0;
                    while (w >= sh.base[slot + 1]) ++slot;
                    idx = w - sh.base[slot];
                    const size_t at = (size_t)slot * p.P + idx;
                    const float4 a = p.in.a[at], b = p.in.b[at];
                    o = v3(a.x, a.y, a.z);
                    d = v3(b.x, b.y, b.z);
                    inv = v3(__fdividef(1.0f, d.x), __fdividef(1.0f, d.y), __fdividef(1.0f, d.z));
                    qr = qray(o, inv, qorg, qcell);
                    t_min = 3.402823466e+38f;
                    hit = -1;
                    best.t = -1.f; best.outside = true; best.aux = v3(0.f, 0.f, 0.f);
                    sp = 0; nc = 0; node = 0;
                    // a NaN direction (glm::refract past the critical angle) hits nothing in the reference
                    walking = d.x == d.x && d.y == d.y && d.z == d.z;
                    has = true;
                }
            }
            if (base + n >= total) exhausted = true;
            idle = __ballot_sync(FULL_MASK, !has);
        }
        if (idle == FULL_MASK) break;                 // only once exhausted: nothing in flight, nothing to fetch

        // ---- walk: box tests only, candidates listed ----
        // The tree is built down to ONE primitive per leaf (PT_GBVH_LEAF == 1), so a child reference is either an inner
        // node or a primitive whose own padded box IS the child box just tested: every step is "fetch one 64-byte node,
        // test its two boxes", whatever the lane stands at, and what follows is a handful of selects -- a primitive child
        // that was hit goes onto the candidate list, a node child is visited (near one first, far one pushed).  (With
        // four primitives per leaf the lanes at leaves ran their own code at 8-11 of 32 threads, a third of the walk's
        // instructions, profiles/r2e_config5_k_isect_dyn_raw.csv.)  The top of the visit stack lives in a register:
        // a pop hands it over at once and the load of the entry below it has the whole next node visit to arrive (the
        // pop's local-memory load used to be the largest single stall of the kernel, 22 % of all samples).
        for (int step = 0; step < PT_DYN_STEPS; ++step) {
            const bool go = has && walking && nc <= PT_GCAND - 2;
            if (!__any_sync(FULL_MASK, go)) break;
            if (go) {
                const float t_lim = fmaf(t_min, 1.0001f, 1e-4f);
                // 32-byte node: two 16-byte loads instead of four (this walk is bound by the L1 load path: every 16-byte load
                // of 32 lanes standing at 32 different nodes is 32 tag look-ups; ncu l1tex throughput 92 % of peak with the
                // 64-byte nodes, profiles/r2e_config5_k_isect_dyn_raw.csv)
                uint4 A, B;
#if PT_NODE_SPLIT
                if (PT_DYN_GNODES > 0 && node < n_staged) {
                    const unsigned at = snodes_s + 32u * (unsigned)node;
                    A = lds_u4(at); B = lds_u4(at + 16u);
                } else {
                    const uint4* q = p.gb.qnodes + 2 * (size_t)node;
                    A = __ldg(q); B = __ldg(q + 1);
                }
#else
                {
                    const uint4* q = node < n_staged ? snodes_q + 2 * node : p.gb.qnodes + 2 * (size_t)node;
                    A = q[0]; B = q[1];
                }
#endif
#if PT_DYN_QFMA
                float t0 = qbox_entry(fmaf(qlo(A.x), qr.ig.x, qr.cg.x), fmaf(qhi(A.y), qr.ig.x, qr.cg.x),
                                      fmaf(qhi(A.x), qr.ig.y, qr.cg.y), fmaf(qlo(A.z), qr.ig.y, qr.cg.y),
                                      fmaf(qlo(A.y), qr.ig.z, qr.cg.z), fmaf(qhi(A.z), qr.ig.z, qr.cg.z), t_lim);
                float t1 = qbox_entry(fmaf(qlo(A.w), qr.ig.x, qr.cg.x), fmaf(qhi(B.x), qr.ig.x, qr.cg.x),
                                      fmaf(qhi(A.w), qr.ig.y, qr.cg.y), fmaf(qlo(B.y), qr.ig.y, qr.cg.y),
                                      fmaf(qlo(B.x), qr.ig.z, qr.cg.z), fmaf(qhi(B.y), qr.ig.z, qr.cg.z), t_lim);
#else
                float4 q0, q1, q2, q3;
                q0.x = qdec_lo(A.x, qcell.x, qorg.x); q0.y = qdec_hi(A.x, qcell.y, qorg.y); q0.z = qdec_lo(A.y, qcell.z, qorg.z);
                q1.x = qdec_hi(A.y, qcell.x, qorg.x); q1.y = qdec_lo(A.z, qcell.y, qorg.y); q1.z = qdec_hi(A.z, qcell.z, qorg.z);
                q2.x = qdec_lo(A.w, qcell.x, qorg.x); q2.y = qdec_hi(A.w, qcell.y, qorg.y); q2.z = qdec_lo(B.x, qcell.z, qorg.z);
                q3.x = qdec_hi(B.x, qcell.x, qorg.x); q3.y = qdec_lo(B.y, qcell.y, qorg.y); q3.z = qdec_hi(B.y, qcell.z, qorg.z);
                float t0 = gbox_entry(q0, q1, o, inv, t_lim), t1 = gbox_entry(q2, q3, o, inv, t_lim);
#endif
                int r0 = (int)B.z, r1 = (int)B.w;
                if (t1 >= 0.0f && (!(t0 >= 0.0f) || t1 < t0)) {            // near child first
                    const int r = r0; r0 = r1; r1 = r;
                    const float t = t0; t0 = t1; t1 = t;
                }
                const bool ok0 = t0 >= 0.0f, ok1 = t1 >= 0.0f;
                if (ok0 && r0 < 0) { cand_prim[nc] = (~r0) >> 3; cand_t[nc] = t0; ++nc; }
                if (ok1 && r1 < 0) { cand_prim[nc] = (~r1) >> 3; cand_t[nc] = t1; ++nc; }
                const bool n0 = ok0 && r0 >= 0, n1 = ok1 && r1 >= 0;
                if (n0 && n1) {                                            // push the far node
                    stack[sp > 0 ? sp - 1 : 0] = top;
                    top = r1;
                    ++sp;
                }
                if (n0 || n1) {
                    node = n0 ? r0 : r1;
                } else if (sp > 0) {                                       // pop
                    node = top;
                    --sp;
                    top = stack[sp > 0 ? sp - 1 : 0];
                } else {
                    walking = false;
                }
            }
        }

        // ---- exact tests of the listed candidates, once some lane needs them ----
        const bool need = has && nc > 0 && (!walking || nc > PT_GCAND - 2);
#if PT_DYN_POOL
        // WARP-COOPERATIVE: the pending candidates of all lanes are pooled and dealt out one per lane, whoever's ray they
        // belong to.  (One ray per lane, a lane testing its own candidates, ran the exact sphere / cube tests -- a third of the
        // kernel's instructions -- at 5 of 32 threads: few lanes have candidates at the same time, and the loop runs as long
        // as the lane with the most.)  A round: every lane proposes its next candidate that is not behind its current hit,
        // nearest first, then -- while the 32 places are not full -- its second, ...; lane i takes pool entry i = (primitive,
        // owner lane), fetches the owner's ray with shuffles and runs the exact test; the results go back to the owner as the
        // lexicographic minimum of (t, scene index) -- a 64-bit shared-memory atomicMin on (float bits, index): t > 0, so the
        // bit patterns order like the values -- which is the hit the reference's ascending loop with its strict `<` keeps
        // (pathtrace.cu:280); the tester that holds the minimum hands over what the normal needs.
        // (with PT_DYN_DRAIN_LANES > 1 the warp waits for that many finished walks -- fuller pools, later pruning -- unless a
        // list is full or no lane can walk on)
        const bool can_walk = has && walking && nc <= PT_GCAND - 2;
        if (__popc(__ballot_sync(FULL_MASK, need)) >= PT_DYN_DRAIN_LANES || __any_sync(FULL_MASK, has && nc > PT_GCAND - 2) ||
            (!__any_sync(FULL_MASK, can_walk) && __any_sync(FULL_MASK, need))) {
            unsigned* pool = wsh.pool[threadIdx.x >> 5];
            unsigned long long* res = wsh.res[threadIdx.x >> 5];
            float4* raux = wsh.aux[threadIdx.x >> 5];
            int k = 0;
            for (;;) {
                int cnt = 0;
                for (int pass = 0; pass < PT_GCAND && cnt < 32; ++pass) {
                    const float lim = fmaf(t_min, 1.0001f, 1e-4f);
                    while (has && k < nc && cand_t[k] > lim) ++k;
                    const bool have = has && k < nc;
                    const unsigned m = __ballot_sync(FULL_MASK, have);
                    if (!m) break;
                    const int at = cnt + __popc(m & ((1u << lane) - 1u));
                    if (have && at < 32) { pool[at] = ((unsigned)cand_prim[k] << 5) | (unsigned)lane; ++k; }
                    cnt = min(32, cnt + __popc(m));
                }
                if (cnt == 0) break;
                res[lane] = ~0ull;
                __syncwarp();
                const unsigned e = lane < cnt ? pool[lane] : (unsigned)lane;
                const int owner = (int)(e & 31u);
                const V3 ro = v3(__shfl_sync(FULL_MASK, o.x, owner), __shfl_sync(FULL_MASK, o.y, owner), __shfl_sync(FULL_MASK, o.z, owner));
                const V3 rd = v3(__shfl_sync(FULL_MASK, d.x, owner), __shfl_sync(FULL_MASK, d.y, owner), __shfl_sync(FULL_MASK, d.z, owner));
                Cand c;
                c.t = -1.0f; c.outside = true; c.aux = v3(0.f, 0.f, 0.f);
                unsigned long long key = ~0ull;
                if (lane < cnt) {
                    const PrimRec& g = p.gb.prims[e >> 5];
                    c = prim_test(g, ro, rd);
                    if (c.t > 0.0f) {
                        key = ((unsigned long long)__float_as_uint(c.t) << 32) | (unsigned)g.pad0;
                        atomicMin(&res[owner], key);
                    }
                }
                __syncwarp();
                if (key != ~0ull && res[owner] == key) raux[owner] = make_float4(c.aux.x, c.aux.y, c.aux.z, c.outside ? 1.0f : 0.0f);
                __syncwarp();
                const unsigned long long mine = res[lane];
                if (mine != ~0ull) {
                    const float t = __uint_as_float((unsigned)(mine >> 32));
                    const int gi = (int)(unsigned)mine;
                    if (t_min > t || (t_min == t && gi < hit)) {
                        const float4 a4 = raux[lane];
                        t_min = t;
                        hit = gi;
                        best.t = t; best.aux = v3(a4.x, a4.y, a4.z); best.outside = a4.w != 0.0f;
                    }
                }
                __syncwarp();
            }
            nc = 0;
        }
#else
        if (__any_sync(FULL_MASK, need)) {
            if (has && nc > 0) {
                int k = 0;
                while (true) {
                    const float lim = fmaf(t_min, 1.0001f, 1e-4f);
                    while (k < nc && cand_t[k] > lim) ++k;
                    if (k >= nc) break;
                    const int ci = cand_prim[k];
                    const PrimRec& g = p.gb.prims[ci];
                    ++k;
#if PT_DYN_CULL
                    // the cheap conservative test against the primitive's true shape (closest_small's): the world box of a
                    // sphere or a rotated cube is mostly empty, and t_lo is a lower bound of the distance the exact test reports
                    float t_lo;
                    if (!cull_test(p.gb.crecs[2 * ci], p.gb.crecs[2 * ci + 1], g, o, d, inv, t_lo) || t_lo > t_min) continue;
#endif
                    const Cand c = prim_test_t<true>(g, o, d);
                    const int gi = g.pad0;
                    // lexicographic minimum of (t, index): what the ascending loop with strict `<` finds
                    if (c.t > 0.0f && (t_min > c.t || (t_min == c.t && gi < hit))) {
                        t_min = c.t;
                        hit = gi;
                        best = c;
                    }
                }
                nc = 0;
            }
        }
#endif

        // ---- finished rays: the hit record goes to the ray's own index ----
        if (has && !walking && nc == 0) {
            float ht = -1.0f;
            V3 hn = v3(0.f, 0.f, 0.f);
            int hm = 0;
            if (hit >= 0) {
                const PrimTail g = load_prim_tail(false, 0u, p.prims + hit);      // (four 16-byte loads)
                hn = prim_normal(g.itr, g.type, best);
                ht = t_min;
                hm = g.material;
            }
            const size_t at = (size_t)slot * p.P + idx;
            p.hit_n[at] = make_float4(ht, hn.x, hn.y, hn.z);
            p.hit_k[at] = hm;
            p.hit_uv[at] = make_float2(0.f, 0.f);
            if (p.cache_mode == 1 && slot == 0) {
                p.cache_n[idx] = make_float4(ht, hn.x, hn.y, hn.z);
                p.cache_k[idx] = hm;
                p.cache_uv[idx] = make_float2(0.f, 0.f);
            }
            has = false;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// k_isect_dyn_mesh: the same dynamic ray fetch for small scenes WITH triangle meshes (the room.json shape:
// a handful of spheres / cubes searched by closest_small, then the walk of the mesh BVH, which dominates and
// diverges just as badly: 7 of 32 threads per instruction at bounce 1, profiles/r19_N3_mesh_room_*).
// Per-lane state is bvh_traverse's (visit stack, listed leaves, nearest triangle so far); the walk visits the
// nodes in the reference's order and the listed triangles are tested in the order met, so the "first one
// met wins a tie" rule is untouched.  Trees deep enough for the reference's stack-overflow drop
// (literal_walk) and the useBVHtree = false mode stay with k_isect.
#ifndef PT_DYNM_NODES
#define PT_DYNM_NODES 0          // nodes of the mesh BVH this kernel stages in shared memory: none (256: 733 spp/s, 64 / 32: 738, 0: 755 on the
                                 // mesh room -- L1 serves the top of the tree as well, and the shared memory it does not lose serves the rest)
#endif
#ifndef PT_DYNM_REFILL
#define PT_DYNM_REFILL 12
#endif

__global__ void __launch_bounds__(PT_THREADS) k_isect_dyn_mesh(const __grid_constant__ RenderParams p)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    DynShared& sh = *reinterpret_cast<DynShared*>(smem_raw);
    BvhRec* snodes = reinterpret_cast<BvhRec*>(smem_raw + sizeof(DynShared));
    // (dynamic layout here: closest_small runs once per ray, before a walk that lives on L1 hits -- the 4.6 KB of the static
    //  stage cost this kernel 2.8 %, profiles/r2_ab_variants.txt)
    const int n_stage = p.mesh.n_smem_nodes < PT_DYNM_NODES ? p.mesh.n_smem_nodes : PT_DYNM_NODES;
    PrimRec* sprims = reinterpret_cast<PrimRec*>(snodes + n_stage);
    float4* crecs = reinterpret_cast<float4*>(sprims + p.gb.n_small);
    const MeshScene& M = p.mesh;
    const int lane = threadIdx.x & 31;
    bounce_barriers_init(sh.bar);
    if (threadIdx.x == 0) {
        int run = 0;
        for (int sl = 0; sl < p.batch; ++sl) {
            sh.base[sl] = run;
            run += p.counts[p.bounce * p.batch + sl];
        }
        sh.base[p.batch] = run;
        sh.total = run;
        if (blockIdx.x == 0) atomicAdd(p.segments, (unsigned long long)run);
    }
    const unsigned snodes_s = smem_u32(snodes);
    stage_begin(p, sh.bar, nullptr, 0, snodes, sprims, true, crecs, n_stage);
    __syncthreads();
    mbar_wait(&sh.bar[0], 0u);
    const int total = sh.total;
    int* ticket = p.tickets + p.depth + p.bounce;       // (the shade launch of this bounce draws from tickets[bounce])

    bool has = false, walking = false;
    int slot = 0, idx = 0;
    V3 o = v3(0.f, 0.f, 0.f), d = o, minv = o;
    float t_min = 3.402823466e+38f, t_cap = 3.402823466e+38f;
    int hit = -1, hit_k = -1;
    Cand best;
    best.t = -1.f; best.outside = true; best.aux = v3(0.f, 0.f, 0.f);
    MeshHit mh = no_mesh_hit();
    int to_visit[BVH_MAX_SEARCH_DEPTH];
    int leaf_first[PT_MLEAVES], leaf_count[PT_MLEAVES];
    float leaf_entry[PT_MLEAVES];
    int sp = 0, cur = 0, nl = 0;
    bool exhausted = false;

    for (;;) {
        // ---- refill: new rays first run the sphere / cube search, whose hit caps the mesh walk ----
        unsigned idle = __ballot_sync(FULL_MASK, !has);
        if (!exhausted && __popc(idle) >= PT_DYNM_REFILL) {
            const int n = __popc(idle);
            int base = 0;
            if (lane == 0) base = atomicAdd(ticket, n);
            base = __shfl_sync(FULL_MASK, base, 0);
            if (!has) {
                const int w = base + __popc(idle & ((1u << lane) - 1u));
                if (w < total) {
                    slot = 0;
                    while (w >= sh.base[slot + 1]) ++slot;
                    idx = w - sh.base[slot];
                    const size_t at = (size_t)slot * p.P + idx;
                    const float4 a = p.in.a[at], b = p.in.b[at];
                    o = v3(a.x, a.y, a.z);
                    d = v3(b.x, b.y, b.z);
                    t_min = 3.402823466e+38f;
                    hit = -1; hit_k = -1;
                    best.t = -1.f; best.outside = true; best.aux = v3(0.f, 0.f, 0.f);
                    mh = no_mesh_hit();
                    sp = 0; cur = 0; nl = 0;
                    walking = d.x == d.x && d.y == d.y && d.z == d.z;      // a NaN direction hits nothing
                    if (walking) {
                        closest_small(sprims, crecs, p.gb.n_small, o, d, t_min, hit, hit_k, best);
                        minv = v3(__frcp_rn(d.x), __frcp_rn(d.y), __frcp_rn(d.z));
                        t_cap = p.gb.cap_mesh_walk ? prune_limit(t_min) : 3.402823466e+38f;
                    }
                    has = true;
                }
            }
            if (base + n >= total) exhausted = true;
            idle = __ballot_sync(FULL_MASK, !has);
        }
        if (idle == FULL_MASK) break;

        // ---- walk of the mesh BVH: box tests only, leaves listed ----
        for (int step = 0; step < PT_DYN_STEPS; ++step) {
            const bool go = has && walking && nl < PT_MLEAVES;
            if (!__any_sync(FULL_MASK, go)) break;
            if (go) {
                const float lim = fminf(prune_limit(mh.t), t_cap);
#if PT_NODE_SPLIT
                float4 w0, w1, w2;                     // BvhRec as three 16-byte words: lo | sub_areas, hi | axis, first rchild - -
                if (PT_DYNM_NODES > 0 && cur < n_stage) {
                    const unsigned at = snodes_s + 48u * (unsigned)cur;
                    w0 = lds_f4(at); w1 = lds_f4(at + 16u); w2 = lds_f4(at + 32u);
                } else {
                    const float4* g = reinterpret_cast<const float4*>(M.nodes + cur);
                    w0 = __ldg(g); w1 = __ldg(g + 1); w2 = __ldg(g + 2);
                }
                const float nlo[3] = {w0.x, w0.y, w0.z}, nhi[3] = {w1.x, w1.y, w1.z};
                const int nd_sub = __float_as_int(w0.w), nd_axis = __float_as_int(w1.w);
                const int nd_first = __float_as_int(w2.x), nd_rchild = __float_as_int(w2.y);
#else
                const BvhRec& nd = cur < M.n_smem_nodes ? snodes[cur] : M.nodes[cur];
                const float* nlo = nd.lo; const float* nhi = nd.hi;
                const int nd_sub = nd.sub_areas, nd_axis = nd.axis, nd_first = nd.first, nd_rchild = nd.rchild;
#endif
                float entry;
                bool descend = false;
                if (bbox_hit(nlo, nhi, o, minv, entry) && !(entry > lim)) {
                    if (nd_sub > 0) {
                        leaf_first[nl] = nd_first; leaf_count[nl] = nd_sub; leaf_entry[nl] = entry; ++nl;
                    } else {
                        descend = true;
                        if ((nd_axis == 0 ? d.x : nd_axis == 1 ? d.y : d.z) < 0.0f) { to_visit[sp++] = cur + 1; cur = nd_rchild; }
                        else { to_visit[sp++] = nd_rchild; cur = cur + 1; }
                    }
                }
                if (!descend) {
                    if (sp == 0) walking = false;
                    else cur = to_visit[--sp];
                }
            }
        }

        // ---- the listed triangles, in the order met, once some lane needs them ----
        const bool need = has && nl > 0 && (!walking || nl == PT_MLEAVES);
        if (__any_sync(FULL_MASK, need)) {
            if (has && nl > 0) {
                int k = 0, i = 0;
                while (true) {
                    const float lim2 = fminf(prune_limit(mh.t), t_cap);
                    while (k < nl && (i >= leaf_count[k] || leaf_entry[k] > lim2)) { ++k; i = 0; }
                    if (k >= nl) break;
                    tri_candidate(M, leaf_first[k] + i, o, d, mh);
                    ++i;
                }
                nl = 0;
            }
        }

        // ---- finished rays ----
        if (has && !walking && nl == 0) {
            if (mh.tri >= 0) mh.id = M.tv[mh.tri].id;
            int owner = -1;
            closest_among_mesh_prims(p, mh, t_min, hit, owner);
            float ht = -1.0f, hu = 0.0f, hv = 0.0f;
            V3 hn = v3(0.f, 0.f, 0.f);
            int hm = 0;
            if (hit >= 0) {
                const bool staged = hit_k >= 0 && sprims[hit_k].pad0 == hit;
                const PrimTail g = load_prim_tail(staged, smem_u32(sprims + hit_k), p.prims + hit);
                hit_attributes(M, g, hit, best, mh, owner, hn, hu, hv);
                ht = t_min;
                hm = g.material;
            }
            const size_t at = (size_t)slot * p.P + idx;
            p.hit_n[at] = make_float4(ht, hn.x, hn.y, hn.z);
            p.hit_k[at] = hm;
            p.hit_uv[at] = make_float2(hu, hv);
            if (p.cache_mode == 1 && slot == 0) {
                p.cache_n[idx] = make_float4(ht, hn.x, hn.y, hn.z);
                p.cache_k[idx] = hm;
                p.cache_uv[idx] = make_float2(hu, hv);
            }
            has = false;
        }
    }
}

// digit of the sort key of the element at sorted position j of this pass
__device__ __forceinline__ int sort_digit(const RenderParams& p, int slot, int j, int pass, int& src)
{
    src = pass == 0 ? j : p.perm_in[(size_t)slot * p.P + j];
    const int* keys = p.cache_mode == 2 ? p.cache_k : p.hit_k + (size_t)slot * p.P;
    return (keys[src] >> (8 * pass)) & 255;
}

// hist[(slot * bins + bin) * tiles + tile] = number of elements of `tile` whose digit is `bin`
// (grid-stride over the batch x tiles work items: most tiles are beyond the live range after a few bounces)
__global__ void __launch_bounds__(PT_THREADS) k_sort_hist(const __grid_constant__ RenderParams p, int pass, int bins)
{
    __shared__ SortShared sh;
    const int items = p.batch * p.tiles;
    for (int w = blockIdx.x; w < items; w += gridDim.x) {
        const int slot = w / p.tiles, tile = w - slot * p.tiles;
        const int n_live = p.counts[p.bounce * p.batch + slot];
        int* hist = p.hist + (size_t)slot * bins * p.tiles;
        if (tile * PT_THREADS >= n_live) {               // every tile writes its column: no memset
            for (int b = threadIdx.x; b < bins; b += PT_THREADS) hist[(size_t)b * p.tiles + tile] = 0;
            continue;
        }
        const int j = tile * PT_THREADS + threadIdx.x;
        const bool active = j < n_live;
        int src = 0;
        const int digit = active ? sort_digit(p, slot, j, pass, src) : 0;
        warp_digit_rank(sh, active, digit, bins);
        __syncthreads();
        for (int b = threadIdx.x; b < bins; b += PT_THREADS) {
            int t = 0;
#pragma unroll
            for (int ww = 0; ww < PT_THREADS / 32; ++ww) t += sh.cnt[ww][b];
            hist[(size_t)b * p.tiles + tile] = t;
        }
        __syncthreads();
    }
}

// perm_out[slot][offset(bin, tile) + rank] = source slot; offsets = exclusive scan of hist
__global__ void __launch_bounds__(PT_THREADS) k_sort_scatter(const __grid_constant__ RenderParams p, int pass, int bins)
{
    __shared__ SortShared sh;
    const int items = p.batch * p.tiles;
    for (int w = blockIdx.x; w < items; w += gridDim.x) {
        const int slot = w / p.tiles, tile = w - slot * p.tiles;
        const int n_live = p.counts[p.bounce * p.batch + slot];
        if (tile * PT_THREADS >= n_live) continue;
        const int* offs = p.hist_scanned + (size_t)slot * bins * p.tiles;
        const int base = offs[0];                        // everything in the slots before this one
        const int j = tile * PT_THREADS + threadIdx.x;
        const bool active = j < n_live;
        int src = 0;
        const int digit = active ? sort_digit(p, slot, j, pass, src) : 0;
        const int rank = warp_digit_rank(sh, active, digit, bins);
        __syncthreads();
        if (active) {
            const int warp = threadIdx.x >> 5;
            int before = 0;
            for (int ww = 0; ww < warp; ++ww) before += sh.cnt[ww][digit];
            const int pos = offs[(size_t)digit * p.tiles + tile] - base + before + rank;
            p.perm_out[(size_t)slot * p.P + pos] = src;
        }
        __syncthreads();
    }
}

// shadeMaterials on the sorted order: slot j of the sorted array is source slot perm[j]; the RNG
// seed uses j, exactly as the reference's shading kernel indexes the sorted arrays.
__global__ void __launch_bounds__(PT_THREADS) k_shade(const __grid_constant__ RenderParams p)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    BounceShared& sh = *reinterpret_cast<BounceShared*>(smem_raw);
    const int n_smem_mats = p.n_mats <= PT_SMEM_MATS ? p.n_mats : 0;
    MatRec* smats = reinterpret_cast<MatRec*>(smem_raw + sizeof(BounceShared));
    const int slot = blockIdx.y;
    const int tile = blockIdx.x;
    const int n_live = p.counts[p.bounce * p.batch + slot];
    int* count_out = p.counts + (p.bounce + 1) * p.batch + slot;
    if (tile * PT_THREADS >= n_live) {
        if (tile == 0 && threadIdx.x == 0) *count_out = 0;
        return;
    }
    if (p.cache_mode == 2 && tile == 0 && threadIdx.x == 0) atomicAdd(p.segments, (unsigned long long)n_live);
    bounce_barriers_init(sh.bar);
    stage_begin(p, sh.bar, smats, n_smem_mats, nullptr, nullptr, false);
    const int j = tile * PT_THREADS + threadIdx.x;
    const bool active = j < n_live;
    const int src = active ? (p.perm_in ? p.perm_in[(size_t)slot * p.P + j] : j) : 0;
    const size_t at = (size_t)slot * p.P + src;
    PathState s = load_path(p.in, at);
    HitRec h;
    {
        const float4 hn = p.cache_mode == 2 ? p.cache_n[src] : p.hit_n[at];
        const float2 huv = p.cache_mode == 2 ? p.cache_uv[src] : p.hit_uv[at];
        h.t = hn.x; h.n = v3(hn.y, hn.z, hn.w);
        h.mat = p.cache_mode == 2 ? p.cache_k[src] : (p.hit_k[at] & 0x00ffffff);
        h.u = huv.x; h.v = huv.y;
    }
    mbar_wait(&sh.bar[0], 0u);                       // the materials
    shade_and_compact(p, sh, smats, n_smem_mats, s, h, active, slot, tile, j, n_live, count_out);
}

// ------------------------------------------------------------------------------------------------
// sendImageToPBO (pathtrace.cu:65-86): note the double-precision multiply by 255.0
__global__ void __launch_bounds__(256) k_tonemap(const float* __restrict__ image, int n_pixels, int iter,
                                                 uchar4* __restrict__ out)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n_pixels) return;
    const float fi = (float)iter;
    int c[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int v = (int)((double)__fdiv_rn(image[3 * (size_t)i + k], fi) * 255.0);
        c[k] = v < 0 ? 0 : (v > 255 ? 255 : v);
    }
    out[i] = make_uchar4((unsigned char)c[0], (unsigned char)c[1], (unsigned char)c[2], 0);
}

// ---- per-stage parity probes on the reference's AoS layouts (tier T1) ---------------------------
__device__ __forceinline__ PathState load_segment(const float* __restrict__ seg)
{
    PathState s;
    s.o = v3(seg[0], seg[1], seg[2]);
    s.d = v3(seg[3], seg[4], seg[5]);
    s.c = v3(seg[6], seg[7], seg[8]);
    s.pixel = __float_as_int(seg[9]);
    s.remaining = __float_as_int(seg[10]);
    s.bounces = __float_as_int(seg[11]);
    return s;
}
__device__ __forceinline__ void store_segment(float* __restrict__ seg, const PathState& s)
{
    seg[0] = s.o.x; seg[1] = s.o.y; seg[2] = s.o.z;
    seg[3] = s.d.x; seg[4] = s.d.y; seg[5] = s.d.z;
    seg[6] = s.c.x; seg[7] = s.c.y; seg[8] = s.c.z;
    seg[9] = __int_as_float(s.pixel);
    seg[10] = __int_as_float(s.remaining);
    seg[11] = __int_as_float(s.bounces);
}

__global__ void __launch_bounds__(256) k_debug_raygen(const __grid_constant__ RenderParams p, int iter,
                                                      float* __restrict__ segments)
{
    const int idx = blockIdx.x * 256 + threadIdx.x;
    if (idx >= p.P) return;
    const int y = idx / p.cam.W, x = idx - y * p.cam.W;
    store_segment(segments + 12 * (size_t)idx, make_camera_path(p.cam, p.opt, iter, p.depth, x, y));
}

__global__ void __launch_bounds__(256) k_debug_intersect(const PrimRec* __restrict__ prims, int n_prims, MeshScene mesh,
                                                         int n, const float* __restrict__ segments,
                                                         float* __restrict__ isects)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const PathState s = load_segment(segments + 12 * (size_t)i);
    float t_min = 3.402823466e+38f;
    int hit = -1;
    Cand best;
    best.t = -1.f; best.outside = true; best.aux = v3(0.f, 0.f, 0.f);
    mesh.n_smem_nodes = 0;
    MeshHit mh = mesh.n_nodes > 0 && !mesh.brute ? bvh_traverse(mesh, nullptr, s.o, s.d) : no_mesh_hit();
    MeshHit last = no_mesh_hit();
    int owner = -1;
    closest_in_chunk(prims, n_prims, 0, s.o, s.d, mesh, mh, last, t_min, hit, best, owner);
    float* o = isects + 7 * (size_t)i;           // the reference memsets the record to 0 first
    V3 nn = v3(0.f, 0.f, 0.f);
    int mat = 0;
    float u = 0.f, v = 0.f;
    if (hit >= 0) {
        hit_attributes(mesh, load_prim_tail(false, 0u, prims + hit), hit, best, mh, owner, nn, u, v);
        mat = prims[hit].material;
    }
    o[0] = hit >= 0 ? t_min : -1.0f;
    o[1] = nn.x; o[2] = nn.y; o[3] = nn.z;
    o[4] = __int_as_float(mat);
    o[5] = u; o[6] = v;
}

__global__ void __launch_bounds__(256) k_debug_shade(const MatRec* __restrict__ mats, MeshScene mesh, OptRec opt, int iter,
                                                     int n, const float* __restrict__ isects, float* __restrict__ segments)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    PathState s = load_segment(segments + 12 * (size_t)i);
    const float* is = isects + 7 * (size_t)i;
    const MatRec m = mats[__float_as_int(is[4])];
    const V3 base = (m.texture_id != -1 && is[0] > 0.0f) ? tex_color(mesh, m.texture_id, is[5], is[6])
                                                         : v3(m.color[0], m.color[1], m.color[2]);
    shade_exact(s, is[0], v3(is[1], is[2], is[3]), m, base, iter, i, opt.rr);
    store_segment(segments + 12 * (size_t)i, s);
}

// ------------------------------------------------------------------------------------------------
size_t bounce_smem_bytes(int n_prims, int n_mats, int n_nodes, int n_gnodes, int n_small)
{
    const int n_chunks = (n_prims + PT_CHUNK - 1) / PT_CHUNK;
    const size_t prim_bytes = n_small ? (size_t)0        // (the small-scene records are in static shared memory: SmallStageT)
                              : n_gnodes ? (size_t)n_gnodes * 4 * sizeof(float4)
                                       : (size_t)(n_chunks > 1 ? 2 * PT_CHUNK : n_prims) * sizeof(PrimRec);
    const size_t mat_bytes = (size_t)(n_mats <= PT_SMEM_MATS ? n_mats : 0) * sizeof(MatRec);
    const size_t node_bytes = (size_t)(n_nodes < PT_SMEM_NODES ? n_nodes : PT_SMEM_NODES) * sizeof(BvhRec);
    return sizeof(BounceShared) + mat_bytes + node_bytes + prim_bytes;
}

// shared memory of the persistent bounce kernel: its control block and survivor stash in front of the staged scene
static size_t persist_smem_bytes(int n_prims, int n_mats, int n_nodes, int n_gnodes, int n_small = 0)
{
    return bounce_smem_bytes(n_prims, n_mats, n_nodes, n_gnodes, n_small) - sizeof(BounceShared) + sizeof(PersistShared) + PT_STASH_BYTES;
}

template <typename K>
static cudaError_t allow_smem(K kernel, int bytes)
{
    return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}

cudaError_t configure_kernels()
{
    const int cap = (int)bounce_smem_bytes(2 * PT_CHUNK, PT_SMEM_MATS, PT_SMEM_NODES);
    const int pcap = (int)persist_smem_bytes(PT_CHUNK, PT_SMEM_MATS, PT_SMEM_NODES, 0);
    cudaError_t e = allow_smem(k_bounce_tile<0, 0>, cap);
    if (e == cudaSuccess) e = allow_smem(k_bounce_tile<1, 0>, cap);
    if (e == cudaSuccess) e = allow_smem(k_bounce_tile<0, 1>, cap);
    if (e == cudaSuccess) e = allow_smem(k_bounce_tile<1, 1>, cap);
    if (e == cudaSuccess) e = allow_smem(k_bounce_tile<0, 2>, cap);
    if (e == cudaSuccess) e = allow_smem(k_bounce_tile<1, 2>, cap);
    if (e == cudaSuccess) e = allow_smem(k_bounce<0, 0>, pcap);
    if (e == cudaSuccess) e = allow_smem(k_bounce<1, 0>, pcap);
    if (e == cudaSuccess) e = allow_smem(k_bounce<2, 0>, pcap);
    if (e == cudaSuccess) e = allow_smem(k_bounce<3, 0>, pcap);
    if (e == cudaSuccess) e = allow_smem(k_bounce<0, 1>, pcap);
    if (e == cudaSuccess) e = allow_smem(k_bounce<1, 1>, pcap);
    if (e == cudaSuccess) e = allow_smem(k_bounce<0, 2>, pcap);
    if (e == cudaSuccess) e = allow_smem(k_bounce<1, 2>, pcap);
    if (e == cudaSuccess) e = allow_smem(k_isect_dyn, (int)(sizeof(DynShared) + (size_t)PT_DYN_GNODES * 2 * sizeof(uint4)));
    if (e == cudaSuccess) e = allow_smem(k_isect_dyn_mesh, cap);

    if (e == cudaSuccess) e = allow_smem(k_isect<0, true>, cap);
    if (e == cudaSuccess) e = allow_smem(k_isect<1, true>, cap);
    if (e == cudaSuccess) e = allow_smem(k_isect<2, true>, cap);
    if (e == cudaSuccess) e = allow_smem(k_isect<1, false>, cap);
    if (e == cudaSuccess) e = allow_smem(k_isect<2, false>, cap);
    if (e == cudaSuccess) e = allow_smem(k_shade, cap);
    return e;
}

cudaError_t launch_raygen(const RenderParams& p, cudaStream_t s)
{
    k_raygen<<<dim3(p.tiles, p.batch), PT_THREADS, 0, s>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_batch_init(const RenderParams& p, cudaStream_t s)
{
    k_batch_init<<<1, 64, 0, s>>>(p);
    return cudaGetLastError();
}

bool bounce_is_persistent(const RenderParams& p)
{
    if (p.cache_mode == 2) return true;
    if (p.gb.n_nodes > 0) return p.gb.n_small > 0 && p.mesh.n_nodes == 0;   // tree walks run one tile per CTA (launch_bounce)
    return p.n_prims <= PT_CHUNK;
}

// CTAs of a persistent launch: as many as are resident at once (more would only queue up behind them)
template <typename K>
static cudaError_t persistent_grid(K kernel, LaunchCfg& cfg, size_t smem, int& grid, int threads = PT_THREADS)
{
    if (cfg.sm_count == 0) {
        int dev = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e == cudaSuccess) e = cudaDeviceGetAttribute(&cfg.sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return e;
    }
    int per_sm = 0;
    const cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem);
    if (e != cudaSuccess) return e;
    grid = cfg.sm_count * (per_sm > 0 ? per_sm : 1);
    if (grid < 1) grid = 1;
    return cudaSuccess;
}

template <int CACHE, int GB>
static cudaError_t launch_persistent(const RenderParams& p, LaunchCfg& cfg, size_t smem, cudaStream_t s)
{
    // the occupancy query is a host-side table lookup, but cache it anyway per shared-memory size (per handle)
    if (cfg.bounce_smem[CACHE][GB] != smem + 1) {
        int g = 0;
        const cudaError_t e = persistent_grid(k_bounce<CACHE, GB>, cfg, smem, g, PT_BTHREADS);
        if (e != cudaSuccess) return e;
        cfg.bounce_grid[CACHE][GB] = g;
        cfg.bounce_smem[CACHE][GB] = smem + 1;
    }
    const int items = p.batch * p.ptiles, grid = cfg.bounce_grid[CACHE][GB];
    k_bounce<CACHE, GB><<<grid < items ? grid : items, PT_BTHREADS, smem, s>>>(p);
    return cudaGetLastError();
}

// which closest-hit search a launch uses: 0 = linear scan, 1 = geometry BVH, 2 = small-scene search
static int search_mode(const RenderParams& p) { return p.gb.n_nodes > 0 ? (p.gb.n_small > 0 ? 2 : 1) : 0; }

cudaError_t launch_bounce(const RenderParams& p, LaunchCfg& cfg, cudaStream_t s)
{
    const int mode = search_mode(p);
    // Scenes whose tiles are long and uneven -- a tree walk per ray: the geometry BVH of a many-primitive scene --
    // run one tile per CTA and leave the balancing to the hardware's CTA scheduler: the persistent kernel's CTA
    // holds its SM share for the slowest warp of every tile in turn (config 5: 85.3 vs 77.3 spp/s,
    // profiles/r5_ab_variants.txt).  The short, even tiles of a small scene are where the persistent form pays.
    if (mode == 1 && p.cache_mode != 2) {
        const dim3 grid(p.tiles, p.batch);
        const size_t smem = bounce_smem_bytes(p.n_prims, p.n_mats, p.mesh.n_smem_nodes, p.gb.n_smem_nodes, 0);
        if (p.cache_mode == 1) k_bounce_tile<1, 1><<<grid, PT_THREADS, smem, s>>>(p);
        else k_bounce_tile<0, 1><<<grid, PT_THREADS, smem, s>>>(p);
        return cudaGetLastError();
    }
    // ... and so does a small scene with triangle meshes: the mesh BVH walk dominates its tiles (+4 % on the mesh room)
    if (mode == 2 && p.mesh.n_nodes > 0 && p.cache_mode != 2) {
        const dim3 grid(p.tiles, p.batch);
        const size_t smem = bounce_smem_bytes(p.n_prims, p.n_mats, p.mesh.n_smem_nodes, p.gb.n_smem_nodes, p.gb.n_small);
        if (p.cache_mode == 1) k_bounce_tile<1, 2><<<grid, PT_THREADS, smem, s>>>(p);
        else k_bounce_tile<0, 2><<<grid, PT_THREADS, smem, s>>>(p);
        return cudaGetLastError();
    }
    if (bounce_is_persistent(p)) {
        const size_t smem = persist_smem_bytes(p.n_prims, p.n_mats, p.mesh.n_smem_nodes, p.gb.n_smem_nodes, p.gb.n_small);
        if (p.cache_mode == 2) return launch_persistent<2, 0>(p, cfg, persist_smem_bytes(0, p.n_mats, 0, 0), s);
        if (p.cache_mode == 1)
            return mode == 2 ? launch_persistent<1, 2>(p, cfg, smem, s) : mode == 1 ? launch_persistent<1, 1>(p, cfg, smem, s) : launch_persistent<1, 0>(p, cfg, smem, s);
        return mode == 2 ? launch_persistent<0, 2>(p, cfg, smem, s) : mode == 1 ? launch_persistent<0, 1>(p, cfg, smem, s) : launch_persistent<0, 0>(p, cfg, smem, s);
    }
    // linear scan over more than one chunk of primitives (geom_bvh off): one CTA per tile, chunks streamed by TMA
    const dim3 grid(p.tiles, p.batch);
    const size_t smem = bounce_smem_bytes(p.n_prims, p.n_mats, p.mesh.n_smem_nodes, 0);
    if (p.cache_mode == 1) k_bounce_tile<1, 0><<<grid, PT_THREADS, smem, s>>>(p);
    else k_bounce_tile<0, 0><<<grid, PT_THREADS, smem, s>>>(p);
    return cudaGetLastError();
}

static size_t dyn_smem_bytes(int n_gnodes)
{
    return sizeof(DynShared) + (size_t)(n_gnodes < PT_DYN_GNODES ? n_gnodes : PT_DYN_GNODES) * 2 * sizeof(uint4);
}

cudaError_t launch_isect(const RenderParams& p, LaunchCfg& cfg, cudaStream_t s, bool* hist_done)
{
    if (hist_done) *hist_done = false;
    const int mode = search_mode(p);
#ifndef PT_NO_DYN
    if (mode == 1 && p.mesh.n_nodes == 0 && p.gb.qnodes) {   // geometry BVH proper, no meshes: dynamic ray fetch
        const size_t smem = dyn_smem_bytes(p.gb.n_nodes);
        if (cfg.dyn_grid == 0) {
            const cudaError_t e = persistent_grid(k_isect_dyn, cfg, dyn_smem_bytes(1 << 30), cfg.dyn_grid);
            if (e != cudaSuccess) return e;
        }
        k_isect_dyn<<<cfg.dyn_grid, PT_THREADS, smem, s>>>(p);
        return cudaGetLastError();
    }
#endif
#ifndef PT_NO_DYN_MESH
    if (mode == 2 && p.mesh.n_nodes > 0 && !p.mesh.literal_walk && !p.mesh.brute) {   // small scene + meshes
        const size_t smem = bounce_smem_bytes(p.n_prims, 0, p.mesh.n_smem_nodes < PT_DYNM_NODES ? p.mesh.n_smem_nodes : PT_DYNM_NODES,
                                              p.gb.n_smem_nodes, p.gb.n_small)
                            - sizeof(BounceShared) + sizeof(DynShared) + (size_t)p.gb.n_small * (sizeof(PrimRec) + 2 * sizeof(float4));
        if (cfg.dyn_mesh_smem != smem) {
            // The walk lives on L1 hits: ask for no more shared-memory carve-out than the four resident CTAs need
            // (+ 1 KB each that the system reserves).  The driver's own choice is larger: mesh room 661 -> 693 spp/s
            // with the 25 % this comes to there.  (For k_isect_dyn, whose CTAs use 1.3 KB, the same hint LOSES 7 %.)
            int dev = 0, sm_smem = 0;
            cudaError_t e = cudaGetDevice(&dev);
            if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sm_smem, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev);
            if (e != cudaSuccess) return e;
            int pct = sm_smem > 0 ? (int)((4 * (smem + 1024) * 100 + (size_t)sm_smem - 1) / (size_t)sm_smem) : 100;
            if (pct > 100) pct = 100;
            e = cudaFuncSetAttribute(k_isect_dyn_mesh, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
            if (e != cudaSuccess) return e;
            e = persistent_grid(k_isect_dyn_mesh, cfg, smem, cfg.dyn_mesh_grid);
            if (e != cudaSuccess) return e;
            cfg.dyn_mesh_smem = smem;
        }
        k_isect_dyn_mesh<<<cfg.dyn_mesh_grid, PT_THREADS, smem, s>>>(p);
        return cudaGetLastError();
    }
#endif
    const size_t smem = bounce_smem_bytes(p.n_prims, 0, p.mesh.n_smem_nodes, p.gb.n_smem_nodes, p.gb.n_small);
    const auto grid_for = [&](int per_cta) { return dim3((unsigned)((p.tiles + per_cta - 1) / per_cta), (unsigned)p.batch); };
    // (a scene without triangle meshes runs the instantiations compiled without the mesh walk: see closest_hit_staged)
    const bool mesh = p.mesh.n_nodes > 0 || !PT_NO_MESH_SPECIAL;
    if (mode == 2 && mesh) k_isect<2, true><<<grid_for(isect_tiles_per_cta<2>()), PT_THREADS, smem, s>>>(p);
    else if (mode == 2) k_isect<2, false><<<grid_for(isect_tiles_per_cta<2>()), PT_THREADS, smem, s>>>(p);
    else if (mode == 1 && mesh) k_isect<1, true><<<grid_for(isect_tiles_per_cta<1>()), PT_THREADS, smem, s>>>(p);
    else if (mode == 1) k_isect<1, false><<<grid_for(isect_tiles_per_cta<1>()), PT_THREADS, smem, s>>>(p);
    else k_isect<0, true><<<grid_for(isect_tiles_per_cta<0>()), PT_THREADS, smem, s>>>(p);
    if (hist_done) *hist_done = p.sort_bins > 0;     // k_isect did the counting half of the sort
    return cudaGetLastError();
}

static cudaError_t stride_grid(const RenderParams& p, LaunchCfg& cfg, int per_sm, int& grid)
{
    if (cfg.sm_count == 0) {
        int dev = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e == cudaSuccess) e = cudaDeviceGetAttribute(&cfg.sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return e;
    }
    const int items = p.batch * p.tiles;
    grid = cfg.sm_count * per_sm;
    if (grid > items) grid = items;
    if (grid < 1) grid = 1;
    return cudaSuccess;
}

cudaError_t launch_sort_hist(const RenderParams& p, LaunchCfg& cfg, int pass, int bins, cudaStream_t s)
{
    int grid = 0;
    const cudaError_t e = stride_grid(p, cfg, 8, grid);
    if (e != cudaSuccess) return e;
    k_sort_hist<<<grid, PT_THREADS, 0, s>>>(p, pass, bins);
    return cudaGetLastError();
}

cudaError_t launch_sort_scatter(const RenderParams& p, LaunchCfg& cfg, int pass, int bins, bool packed, cudaStream_t s)
{
    int grid = 0;
    const cudaError_t e = stride_grid(p, cfg, 8, grid);
    if (e != cudaSuccess) return e;
    if (packed) k_sort_scatter_packed<<<grid, PT_THREADS, 0, s>>>(p, bins);
    else k_sort_scatter<<<grid, PT_THREADS, 0, s>>>(p, pass, bins);
    return cudaGetLastError();
}

cudaError_t launch_shade(const RenderParams& p, LaunchCfg& cfg, cudaStream_t s)
{
#ifdef PT_OLD_SHADE
    k_shade<<<dim3(p.tiles, p.batch), PT_THREADS, bounce_smem_bytes(0, p.n_mats, 0), s>>>(p);
    return cudaGetLastError();
#else
    // the persistent bounce kernel in its shade-only mode: tiles in ticket order, look-back off the critical path
    return launch_persistent<3, 0>(p, cfg, persist_smem_bytes(0, p.n_mats, 0, 0), s);
#endif
}

cudaError_t launch_tonemap(const float* image, int n_pixels, int iter, uchar4* out, cudaStream_t s)
{
    k_tonemap<<<(n_pixels + 255) / 256, 256, 0, s>>>(image, n_pixels, iter, out);
    return cudaGetLastError();
}

cudaError_t launch_debug_raygen(const RenderParams& p, int iter, float* segments, cudaStream_t s)
{
    k_debug_raygen<<<(p.P + 255) / 256, 256, 0, s>>>(p, iter, segments);
    return cudaGetLastError();
}

cudaError_t launch_debug_intersect(const PrimRec* prims, int n_prims, MeshScene mesh, int n, const float* segments,
                                   float* isects, cudaStream_t s)
{
    k_debug_intersect<<<(n + 255) / 256, 256, 0, s>>>(prims, n_prims, mesh, n, segments, isects);
    return cudaGetLastError();
}

cudaError_t launch_debug_shade(const MatRec* mats, MeshScene mesh, OptRec opt, int iter, int n, const float* isects,
                               float* segments, cudaStream_t s)
{
    k_debug_shade<<<(n + 255) / 256, 256, 0, s>>>(mats, mesh, opt, iter, n, isects, segments);
    return cudaGetLastError();
}

}  // namespace ptk
