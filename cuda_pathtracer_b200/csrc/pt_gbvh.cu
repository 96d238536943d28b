// pt_gbvh.cu -- host-side build of the geometry BVH (GeomBvh, pt_kernels.cuh): an index over the
// sphere / cube primitives of a scene.  The reference has no such structure -- computeIntersections
// tests every Geom for every ray (pathtrace.cu:247-290) -- and nothing here changes WHICH hit is
// found: the tree only prunes primitives the ray cannot touch, every candidate still goes through
// the exact (bit-reproducing) intersection routines, and the winner is the minimum of (t, index).
//
// Why the boxes are padded.  The reference decides hit / miss in fp32 in the primitive's object
// space, so it also reports hits for rays that pass just OUTSIDE the exact shape:
//   * sphere (intersections.cu:60-115): radicand = dot(o,d)^2 - (dot(o,o) - r^2) cancels two numbers
//     of size (D/s)^2 (D = distance ray origin -> centre, s = smallest scale), error of a few
//     ulp((D/s)^2), i.e. the sphere appears up to ~ 4e-7 * D^2 / s larger in world space
//     (0.006 for D = 70, s = 0.3);
//   * cube: slab distances carry a few ulp(D/s) in object space, ~ 1e-6 * D in world space.
// A primitive must be found whenever the reference would find it, and the distance the tree code
// prunes with must never exceed the distance the reference would report.  Two margins per primitive:
//   d_hit = 24 * eps * D_max (+ 32 * eps * D_max^2 / s_min for a sphere), eps = 2^-23, D_max = diagonal of
//           everything a ray can start from (all primitives, all meshes, the camera): covers the rounding
//           of the reference's hit / miss decision and of the conservative tests here;
//   d_t   = 1.05e-4 * s_max + 24 * eps * D_max: the reference reports the distance to a point pulled back
//           by 1e-4 OBJECT units along the ray (getPointOnRay, intersections.h:29-32), up to 1e-4 * s_max
//           in world units.
// The tree boxes are grown by d_hit + d_t (they decide hit / miss and bound the distance at once).  The
// per-primitive cull records of small scenes (crecs, see closest_small in pt_kernels.cu) keep the two
// apart: a ray that has just left a surface starts 1e-4 outside it -- inside the d_t margin but outside
// the d_hit one -- and must not cost an exact test of the surface it leaves.
// tests/test_pt_gpu.py checks BVH on / off bit-identity on every scene, and
// the 600-primitive golden render of the reference's CUDA build pins the BVH path directly.
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <vector>

#include "pt_kernels.cuh"

namespace ptk {

namespace {

struct Box3 {
    double lo[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, hi[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
    void grow(const double p[3]) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], p[a]); hi[a] = std::max(hi[a], p[a]); } }
    void grow(const Box3& b) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], b.lo[a]); hi[a] = std::max(hi[a], b.hi[a]); } }
    double area() const
    {
        const double d[3] = {hi[0] - lo[0], hi[1] - lo[1], hi[2] - lo[2]};
        if (d[0] < 0) return 0.0;
        return 2.0 * (d[0] * d[1] + d[0] * d[2] + d[1] * d[2]);
    }
};

struct Ref {
    int prim;
    int ex;                        // position in the per-primitive side table (refs get reordered by the build)
    Box3 box;
    double c[3];
};

struct Tmp {                       // build-time node
    Box3 box;
    int left = -1, right = -1;     // children (Tmp indices) or -1
    int first = 0, count = 0;      // leaf range in the ordered refs
};

constexpr int kBins = 16;

int build(std::vector<Tmp>& nodes, std::vector<Ref>& refs, int begin, int end, int depth)
{
    const int me = (int)nodes.size();
    nodes.emplace_back();
    Box3 box, cbox;
    for (int i = begin; i < end; ++i) { box.grow(refs[i].box); cbox.grow(refs[i].c); }
    nodes[me].box = box;
    const int n = end - begin;
    if (n <= PT_GBVH_LEAF) { nodes[me].first = begin; nodes[me].count = n; return me; }

    int axis = 0;
    for (int a = 1; a < 3; ++a) if (cbox.hi[a] - cbox.lo[a] > cbox.hi[axis] - cbox.lo[axis]) axis = a;
    const double ext = cbox.hi[axis] - cbox.lo[axis];
    int mid = -1;
    if (ext > 0.0 && depth < 32) {                              // binned surface-area heuristic
        Box3 bb[kBins];
        int bc[kBins] = {0};
        auto bin_of = [&](const Ref& r) { return std::min(kBins - 1, (int)(kBins * (r.c[axis] - cbox.lo[axis]) / ext)); };
        for (int i = begin; i < end; ++i) { const int b = bin_of(refs[i]); bc[b]++; bb[b].grow(refs[i].box); }
        double best = DBL_MAX;
        int split = -1;
        for (int s = 1; s < kBins; ++s) {
            Box3 l, r;
            int nl = 0, nr = 0;
            for (int b = 0; b < s; ++b) if (bc[b]) { l.grow(bb[b]); nl += bc[b]; }
            for (int b = s; b < kBins; ++b) if (bc[b]) { r.grow(bb[b]); nr += bc[b]; }
            if (!nl || !nr) continue;
            const double cost = nl * l.area() + nr * r.area();
            if (cost < best) { best = cost; split = s; }
        }
        if (split > 0)
            mid = (int)(std::partition(refs.begin() + begin, refs.begin() + end, [&](const Ref& r) { return bin_of(r) < split; }) -
                        refs.begin());
    }
    if (mid <= begin || mid >= end) {                           // equal centroids / depth guard: median by index
        mid = begin + n / 2;
        if (ext > 0.0)
            std::nth_element(refs.begin() + begin, refs.begin() + mid, refs.begin() + end,
                             [&](const Ref& a, const Ref& b) { return a.c[axis] < b.c[axis]; });
    }
    const int l = build(nodes, refs, begin, mid, depth + 1);
    const int r = build(nodes, refs, mid, end, depth + 1);
    nodes[me].left = l;
    nodes[me].right = r;
    return me;
}

}  // namespace

void build_geom_bvh(const std::vector<PrimRec>& prims, const double cam_pos[3], const double mesh_lo[3], const double mesh_hi[3],
                    bool have_mesh_bounds, std::vector<float4>& nodes_out, std::vector<PrimRec>& prims_out,
                    std::vector<float4>& boxes_out, std::vector<int>& mesh_prims_out, std::vector<uint4>& qnodes_out,
                    float qorg[3], float qcell[3])
{
    nodes_out.clear(); prims_out.clear(); boxes_out.clear(); mesh_prims_out.clear(); qnodes_out.clear();
    struct Extra { double s[3]; double d_hit, d_t; bool aligned, finite; Box3 tight; };
    std::vector<Extra> extra;
    const double eps = 1.1920928955078125e-7;
    std::vector<Ref> refs;
    std::vector<double> smin;
    Box3 world;
    world.grow(cam_pos);
    if (have_mesh_bounds) { world.grow(mesh_lo); world.grow(mesh_hi); }
    for (int i = 0; i < (int)prims.size(); ++i) {
        const PrimRec& g = prims[i];
        if (g.type == 2) { mesh_prims_out.push_back(i); continue; }
        Ref r;
        r.prim = i;
        for (int k = 0; k < 8; ++k) {                            // the unit cube bounds the r = 0.5 sphere too
            const double u[3] = {k & 1 ? 0.5 : -0.5, k & 2 ? 0.5 : -0.5, k & 4 ? 0.5 : -0.5};
            double w[3];
            for (int row = 0; row < 3; ++row)
                w[row] = (double)g.fwd[0 * 3 + row] * u[0] + (double)g.fwd[1 * 3 + row] * u[1] + (double)g.fwd[2 * 3 + row] * u[2] +
                         (double)g.fwd[3 * 3 + row];
            r.box.grow(w);
        }
        Extra ex{};
        ex.finite = true;
        for (int k = 0; k < 12; ++k) ex.finite = ex.finite && std::isfinite(g.fwd[k]) && std::isfinite(g.inv[k]);
        double s = DBL_MAX;                                      // column norms of the linear part = the three scales
        for (int c = 0; c < 3; ++c) {
            ex.s[c] = std::sqrt((double)g.fwd[c * 3] * g.fwd[c * 3] + (double)g.fwd[c * 3 + 1] * g.fwd[c * 3 + 1] +
                                (double)g.fwd[c * 3 + 2] * g.fwd[c * 3 + 2]);
            s = std::min(s, ex.s[c]);
        }
        // axis aligned: every object axis maps onto one world axis -- up to 1e-5 of its length, so that a
        // 90-degree rotation (cos = -4.4e-8 in fp32) still counts; the world box of the corners then exceeds
        // the cube by that little, which a conservative test may
        ex.aligned = true;
        for (int c = 0; c < 3; ++c) {
            int big = 0;
            for (int row = 0; row < 3; ++row) big += std::fabs((double)g.fwd[c * 3 + row]) > 1e-5 * ex.s[c];
            if (big != 1) ex.aligned = false;
        }
        for (int row = 0; row < 3; ++row) {
            int big = 0;
            for (int c = 0; c < 3; ++c) big += std::fabs((double)g.fwd[c * 3 + row]) > 1e-5 * ex.s[c];
            if (big != 1) ex.aligned = false;
        }
        ex.tight = r.box;
        r.ex = (int)extra.size();
        extra.push_back(ex);
        smin.push_back(s);
        world.grow(r.box);
        refs.push_back(r);
    }
    if (refs.empty()) return;
    double diag2 = 0.0;
    for (int a = 0; a < 3; ++a) diag2 += (world.hi[a] - world.lo[a]) * (world.hi[a] - world.lo[a]);
    const double dmax = std::sqrt(diag2) * 1.05 + 1.0;           // + lens offsets / the 1e-4 surface offsets
    for (size_t i = 0; i < refs.size(); ++i) {
        Extra& ex = extra[i];
        ex.d_hit = 24.0 * eps * dmax;
        if (prims[(size_t)refs[i].prim].type == 0) ex.d_hit += smin[i] > 0.0 ? 32.0 * eps * dmax * dmax / smin[i] : dmax;   // sphere
        ex.d_t = 1.05e-4 * std::max(ex.s[0], std::max(ex.s[1], ex.s[2])) + 24.0 * eps * dmax;
        double pad = ex.d_hit + ex.d_t;
        if (!(pad < dmax)) pad = dmax;                           // degenerate scale: the box covers everything
        for (int a = 0; a < 3; ++a) {
            if (!std::isfinite(refs[i].box.lo[a]) || !std::isfinite(refs[i].box.hi[a])) { refs[i].box.lo[a] = -1e30; refs[i].box.hi[a] = 1e30; ex.finite = false; }
            refs[i].box.lo[a] -= pad;
            refs[i].box.hi[a] += pad;
            refs[i].c[a] = 0.5 * (refs[i].box.lo[a] + refs[i].box.hi[a]);
        }
    }
    std::vector<Tmp> tmp;
    tmp.reserve(2 * refs.size());
    build(tmp, refs, 0, (int)refs.size(), 0);

    prims_out.resize(refs.size());
    for (size_t i = 0; i < refs.size(); ++i) {
        prims_out[i] = prims[(size_t)refs[i].prim];
        prims_out[i].pad0 = refs[i].prim;                        // index in the scene's primitive array
    }
    // inner nodes in breadth-first order (the top levels first: they are the ones staged in shared memory)
    std::vector<int> order, id_of(tmp.size(), -1);
    if (tmp[0].left < 0) {                                       // a single leaf: give it a parent with two equal children
        Tmp root;
        root.box = tmp[0].box;
        root.left = root.right = 0;
        tmp.push_back(root);
        order.push_back((int)tmp.size() - 1);
        id_of.push_back(0);
    } else {
        order.push_back(0);
        id_of[0] = 0;
    }
    for (size_t at = 0; at < order.size(); ++at) {
        const Tmp& t = tmp[(size_t)order[at]];
        for (int child : {t.left, t.right})
            if (tmp[(size_t)child].left >= 0 && id_of[(size_t)child] < 0) { id_of[(size_t)child] = (int)order.size(); order.push_back(child); }
    }
    auto ref_of = [&](int child) {
        const Tmp& c = tmp[(size_t)child];
        return c.left >= 0 ? id_of[(size_t)child] : ~((c.first << 3) | (c.count - 1));
    };
    auto lo4 = [](const Box3& b, int w) {                        // round outwards
        float4 q;
        q.x = std::nextafterf((float)b.lo[0], -FLT_MAX); q.y = std::nextafterf((float)b.lo[1], -FLT_MAX);
        q.z = std::nextafterf((float)b.lo[2], -FLT_MAX); std::memcpy(&q.w, &w, 4);
        return q;
    };
    auto hi4 = [](const Box3& b, int w) {
        float4 q;
        q.x = std::nextafterf((float)b.hi[0], FLT_MAX); q.y = std::nextafterf((float)b.hi[1], FLT_MAX);
        q.z = std::nextafterf((float)b.hi[2], FLT_MAX); std::memcpy(&q.w, &w, 4);
        return q;
    };
    boxes_out.resize(4 * refs.size());
    for (size_t i = 0; i < refs.size(); ++i) {
        boxes_out[2 * i + 0] = lo4(refs[i].box, 0);
        boxes_out[2 * i + 1] = hi4(refs[i].box, 0);
    }
    // Cull records of the small-scene path, two float4 per primitive, in the same (leaf) order, behind the boxes:
    //   kind 0  axis-aligned cube: {world lo - d_hit, kind} {world hi + d_hit, d_t}
    //   kind 1  cube:              {object half extents 0.5 + d_hit / s_a, kind} {d_t / s_a, -}
    //   kind 2  sphere:            {object-space error of a transformed point, d_t / s_min, -, kind}
    //   kind 3  degenerate transform: always a candidate
    auto up = [](double v) { return std::nextafterf((float)v, FLT_MAX); };
    for (size_t i = 0; i < refs.size(); ++i) {
        const Extra& ex = extra[(size_t)refs[i].ex];
        const int type = prims[(size_t)refs[i].prim].type;
        const double smin_i = std::min(ex.s[0], std::min(ex.s[1], ex.s[2]));
        int kind = 3;
        float4 q0 = {0, 0, 0, 0}, q1 = {0, 0, 0, 0};
        if (ex.finite && smin_i > 0.0 && ex.d_hit + ex.d_t < dmax) {
            if (type == 1 && ex.aligned) {
                kind = 0;
                Box3 b = ex.tight;
                for (int a = 0; a < 3; ++a) { b.lo[a] -= ex.d_hit; b.hi[a] += ex.d_hit; }
                q0 = lo4(b, 0);
                q1 = hi4(b, 0);
                q1.w = up(ex.d_t);
            } else if (type == 1) {
                kind = 1;
                q0.x = up(0.5 + ex.d_hit / ex.s[0]); q0.y = up(0.5 + ex.d_hit / ex.s[1]); q0.z = up(0.5 + ex.d_hit / ex.s[2]);
                q1.x = up(ex.d_t / ex.s[0]); q1.y = up(ex.d_t / ex.s[1]); q1.z = up(ex.d_t / ex.s[2]);
            } else if (type == 0) {
                // object-space error of the transformed origin (the reference's rounding and this code's) and
                // the distance margin; the radius margin depends on where the ray starts (closest_small)
                kind = 2;
                q0.x = up(8.0 * eps * dmax / smin_i);
                q0.y = up(ex.d_t / smin_i);
            }
        }
        std::memcpy(&q0.w, &kind, 4);
        boxes_out[2 * refs.size() + 2 * i + 0] = q0;
        boxes_out[2 * refs.size() + 2 * i + 1] = q1;
    }
    nodes_out.resize(4 * order.size());
    for (size_t i = 0; i < order.size(); ++i) {
        const Tmp& t = tmp[(size_t)order[i]];
        const Tmp &l = tmp[(size_t)t.left], &r = tmp[(size_t)t.right];
        nodes_out[4 * i + 0] = lo4(l.box, ref_of(t.left));
        nodes_out[4 * i + 1] = hi4(l.box, ref_of(t.right));
        nodes_out[4 * i + 2] = lo4(r.box, 0);
        nodes_out[4 * i + 3] = hi4(r.box, 0);
    }

    // The same tree in 32-byte nodes for k_isect_dyn, whose walk is bound by the L1 load path (one 16-byte load per lane
    // is one tag look-up per lane: four per 64-byte node): 16-bit coordinates on one grid over the root box, lower
    // corners rounded down and upper corners rounded up by one extra cell -- the boxes only ever grow (a cell is 1/65000
    // of the scene: 1e-3 for config 5's 80 units), and the decoded value q * cell + org is off by an ulp at most, which the
    // extra cell covers.  Skipped (qnodes_out stays empty) when a box is not finite.
    bool finite = true;
    for (const Extra& ex : extra) finite = finite && ex.finite;
    const Box3& root = tmp[(size_t)order[0]].box;
    for (int a = 0; a < 3; ++a) finite = finite && std::isfinite(root.lo[a]) && std::isfinite(root.hi[a]) && root.hi[a] - root.lo[a] < 1e30;
    if (!finite) return;
    double org[3], cell[3];
    for (int a = 0; a < 3; ++a) {
        const double ext = std::max(root.hi[a] - root.lo[a], 1e-6);
        cell[a] = ext / 65000.0;
        org[a] = root.lo[a] - 200.0 * cell[a];               // 65000 cells of the scene + 200 spare ones on either side
        qorg[a] = (float)org[a];
        qcell[a] = (float)cell[a];
    }
    auto q_lo = [&](double v, int a) {                            // against the FLOAT grid the kernel decodes with
        const double q = std::floor((v - (double)qorg[a]) / (double)qcell[a]) - 1.0;
        return (unsigned)std::min(65535.0, std::max(0.0, q));
    };
    auto q_hi = [&](double v, int a) {
        const double q = std::ceil((v - (double)qorg[a]) / (double)qcell[a]) + 1.0;
        return (unsigned)std::min(65535.0, std::max(0.0, q));
    };
    qnodes_out.resize(2 * order.size());
    for (size_t i = 0; i < order.size(); ++i) {
        const Tmp& t = tmp[(size_t)order[i]];
        const Box3 &l = tmp[(size_t)t.left].box, &r = tmp[(size_t)t.right].box;
        const unsigned c[12] = {q_lo(l.lo[0], 0), q_lo(l.lo[1], 1), q_lo(l.lo[2], 2), q_hi(l.hi[0], 0), q_hi(l.hi[1], 1), q_hi(l.hi[2], 2),
                                q_lo(r.lo[0], 0), q_lo(r.lo[1], 1), q_lo(r.lo[2], 2), q_hi(r.hi[0], 0), q_hi(r.hi[1], 1), q_hi(r.hi[2], 2)};
        uint4 A, B;
        A.x = c[0] | (c[1] << 16); A.y = c[2] | (c[3] << 16); A.z = c[4] | (c[5] << 16); A.w = c[6] | (c[7] << 16);
        B.x = c[8] | (c[9] << 16); B.y = c[10] | (c[11] << 16);
        B.z = (unsigned)ref_of(t.left); B.w = (unsigned)ref_of(t.right);
        qnodes_out[2 * i] = A;
        qnodes_out[2 * i + 1] = B;
    }
}

}  // namespace ptk
