// pt_kernels.cuh -- launcher interface between pt_kernels.cu (device code) and pt_api.cu (the
// C ABI of include/b200_pt.h).  Not part of the public boundary.
#pragma once
#include <cuda_runtime.h>
#include <vector>
#include "pt_shade.cuh"

namespace ptk {

constexpr int PT_THREADS = 256;            // paths per tile / threads per CTA
#ifndef PT_PWARPS_N
#define PT_PWARPS_N 7
#endif
constexpr int PT_PWARPS = PT_PWARPS_N;      // compute warps of the persistent fused bounce kernel; one more warp
                                           // only runs the look-backs of the CTA's tiles (k_bounce, pt_kernels.cu)
constexpr int PT_PTILE = PT_PWARPS * 32;   // paths per tile of that kernel
constexpr int PT_BTHREADS = PT_PTILE + 32; // threads of its CTAs
constexpr int PT_MAX_BATCH = 64;           // iterations in flight (b200pt_options::batch upper bound)
constexpr int PT_CHUNK = 256;              // primitives staged per shared-memory buffer (40 KB)
constexpr int PT_SMEM_MATS = 256;          // materials kept in shared memory (12 KB)
constexpr int PT_SMEM_NODES = 256;         // BVH nodes kept in shared memory (12 KB): DFS order => the top of the tree
#ifndef PT_SMEM_GNODES_N
#define PT_SMEM_GNODES_N 32
#endif
constexpr int PT_SMEM_GNODES = PT_SMEM_GNODES_N;   // geometry-BVH nodes staged in shared memory (2 KB): breadth-first order => the top
                                           // 5 levels, which every ray visits.  More costs more than it saves now that these scenes run
                                           // one tile per CTA (every tile re-stages them): config 5 89.4 spp/s with 32, 87.2 with 64,
                                           // 85.2 with 192, 83.5 with 512, 63.2 with 1024 (profiles/r5_ab_variants.txt); the rest of
                                           // the tree is served by L1 / L2
#ifndef PT_GBVH_LEAF_N
#define PT_GBVH_LEAF_N 1
#endif
constexpr int PT_GBVH_LEAF = PT_GBVH_LEAF_N;   // primitives per geometry-BVH leaf (at most 8: 3 bits of the child reference).  1: a child
                                           // reference is an inner node or ONE primitive whose padded box is the child box itself
                                           // (what k_isect_dyn's uniform walk step needs; 4 was the round-1 value)
constexpr int PT_GBVH_AUTO = 1;            // geom_bvh = auto: use the BVH from this many sphere / cube primitives (with fewer
                                           // than a leaf's worth the tree is one node and the per-primitive boxes do the culling)
constexpr int PT_SMALL_MAX = 24;           // sphere / cube primitives up to which the small-scene search replaces the tree
constexpr int PT_DESC_STRIDE = 16;         // 64-bit words between look-back descriptors: one 128-B line each
                                           // (packed descriptors make the lines around the prefix frontier an
                                           // L2 hot spot, see lookback.cuh)

// SoA path pool: three 16-byte words per path.
//   a = origin.xyz    | pixelIndex
//   b = direction.xyz | remainingBounces
//   c = colour.xyz    | bounces
struct Pool {
    float4* a;
    float4* b;
    float4* c;
};

// Geometry BVH: an index over the sphere / cube primitives that replaces the linear scan of
// computeIntersections (pathtrace.cu:247-290) when the scene has many of them.  It only PRUNES:
// every primitive whose (padded, see pt_api.cu) world box the ray may touch still goes through
// the exact prim_test, and the winner is the lexicographic minimum of (t, index) -- the same hit
// the reference's ascending loop with its strict `<` selects.
//   node = 4 x float4: {lo_L, ref_L} {hi_L, ref_R} {lo_R, -} {hi_R, -}
//   ref >= 0: inner node index; ref < 0: leaf, v = ~ref, first = v >> 3, count = (v & 7) + 1
struct GeomBvh {
    const float4* nodes;
    const PrimRec* prims;              // copies in leaf order; pad0 = index in the scene's primitive array
    const float4* boxes;               // padded world box of each of them: {lo, -} {hi, -}
    const float4* crecs;               // cull records (two float4 per primitive, same order; pt_gbvh.cu)
    const uint4* qnodes;               // the same tree in 32-byte nodes (two uint4): both child boxes as 12 unsigned 16-bit
                                       // coordinates on ONE grid over the scene (qorg + q * qcell, rounded outwards), then
                                       // the two child references.  nullptr: not available (a box is not finite).
                                       //   A = {c0.lo.x | c0.lo.y << 16, c0.lo.z | c0.hi.x << 16, c0.hi.y | c0.hi.z << 16, c1.lo.x | c1.lo.y << 16}
                                       //   B = {c1.lo.z | c1.hi.x << 16, c1.hi.y | c1.hi.z << 16, ref0, ref1}
    float qorg[3], qcell[3];
    const int* mesh_prims;             // indices of the MESH primitives, ascending
    int n_nodes;                       // 0 = not in use (linear scan over the staged primitives)
    int n_smem_nodes;
    int n_mesh_prims;
    int n_small;                       // > 0: so few sphere / cube primitives (<= PT_SMALL_MAX) that the tree is
                                       // skipped: all of them and their cull records are staged in shared memory
    int cap_mesh_walk;                 // no sphere / cube has a textured material: a mesh hit behind the nearest
                                       // sphere / cube hit cannot show anywhere (not even as the stale texture
                                       // coordinate of pathtrace.cu:267,283), so the mesh walk may stop there
};

struct RenderParams {
    Pool in, out;                      // ping-pong pools, batch slot s at offset s * P
    int* counts;                       // [(depth + 1)][batch]: live paths entering each bounce
    unsigned long long* desc;          // [batch][tiles]: look-back descriptors (epoch-tagged)
    unsigned long long* segments;      // running sum of live paths entering intersection
    float* image;                      // float SUM image, 3 floats per pixel
    const PrimRec* prims;
    const MatRec* mats;
    int n_prims, n_mats;
    CamRec cam;
    OptRec opt;
    int depth;
    int P;                             // pixels = paths per iteration
    int tiles;                         // ceil(P / PT_THREADS)
    int ptiles;                        // ceil(P / PT_PTILE): tiles per slot of the persistent bounce kernel
    int* tickets;                      // [2 * depth]: next work item of each bounce's persistent launches (zeroed by k_raygen):
                                       // [bounce] shading / fused bounce, [depth + bounce] closest hit with dynamic ray fetch
    int batch;                         // iterations in flight in this launch (grid.y)
    int iter0, stride;                 // slot s traces iteration iter0 + s * stride
    int bounce;                        // which bounce this launch is
    unsigned epoch;                    // look-back epoch of this launch
    // split pipeline (material sort) and first-bounce cache
    float4* hit_n;                     // [batch][P] hit record {t, normal}
    int* hit_k;                        // [batch][P] hit record material id (the sort key; 0 on a miss)
    const int* perm_in;                // [batch][P] sorted position -> source slot (nullptr = identity)
    int perm_shared;                   // 1: perm_in is ONE [P] permutation shared by all slots (the cached bounce-0 order)
    int sort_bins;                     // > 0: the closest-hit kernel also writes this launch's digit histogram (pass 0)
    int* perm_out;
    int* hist;                         // [batch][bins][tiles] digit counts per tile
    const int* hist_scanned;           // exclusive scan of hist (flattened)
    float4* cache_n;                   // [P] first-bounce cache
    int* cache_k;
    int cache_mode;                    // at bounce 0: 0 = off, 1 = fill, 2 = use
    int raygen;                        // 1 (bounce 0 of the persistent fused kernel only): the paths of this launch are the camera
                                       // paths, generated in the kernel instead of loaded (launch_batch_init replaces launch_raygen)
    float2* hit_uv;                    // [batch][P] hit record texture coordinate (mesh hits)
    float2* cache_uv;                  // [P]
    MeshScene mesh;                    // triangle meshes / BVH / textures (n_nodes == 0: none)
    GeomBvh gb;
};

// Launch configuration of ONE renderer handle: SM count of its device and the resident-CTA grids of the
// persistent kernels, computed on first use for the device the handle was created on.  Kept per handle (not in
// function-local statics) so that renderers on different GPUs, or driven from different host threads, of one
// process do not share or race on it (include/b200_pt.h: several renderers may coexist).
struct LaunchCfg {
    int sm_count = 0;
    size_t bounce_smem[4][3] = {};     // [CACHE][GB]: shared memory the cached grid was computed for (0 = not yet)
    int bounce_grid[4][3] = {};
    int dyn_grid = 0;                  // k_isect_dyn
    int dyn_mesh_grid = 0;             // k_isect_dyn_mesh
    size_t dyn_mesh_smem = ~(size_t)0;
};

size_t bounce_smem_bytes(int n_prims, int n_mats, int n_nodes, int n_gnodes = 0, int n_small = 0);
cudaError_t launch_raygen(const RenderParams& p, cudaStream_t s);
// the bookkeeping half of k_raygen alone (live counts = P, ticket counters = 0) for a batch whose bounce 0 generates its own rays
cudaError_t launch_batch_init(const RenderParams& p, cudaStream_t s);
cudaError_t launch_bounce(const RenderParams& p, LaunchCfg& cfg, cudaStream_t s);   // fused (honours p.cache_mode)
// split pipeline: closest hit -> hit records; one radix pass = hist, (scan by the caller), scatter; shade
// hist_done (may be null): set when the launch also wrote the digit histogram of sort pass 0 (p.sort_bins > 0)
cudaError_t launch_isect(const RenderParams& p, LaunchCfg& cfg, cudaStream_t s, bool* hist_done = nullptr);
cudaError_t launch_sort_hist(const RenderParams& p, LaunchCfg& cfg, int pass, int bins, cudaStream_t s);
// packed: the keys carry their in-tile rank (written by a closest-hit launch that reported hist_done)
cudaError_t launch_sort_scatter(const RenderParams& p, LaunchCfg& cfg, int pass, int bins, bool packed, cudaStream_t s);
cudaError_t launch_shade(const RenderParams& p, LaunchCfg& cfg, cudaStream_t s);
cudaError_t launch_tonemap(const float* image, int n_pixels, int iter, uchar4* out, cudaStream_t s);
// per-stage probes on the reference's AoS layouts (PathSegment 12 words, ShadeableIntersection 7 words)
cudaError_t launch_debug_raygen(const RenderParams& p, int iter, float* segments, cudaStream_t s);
cudaError_t launch_debug_intersect(const PrimRec* prims, int n_prims, MeshScene mesh, int n, const float* segments,
                                   float* isects, cudaStream_t s);
cudaError_t launch_debug_shade(const MatRec* mats, MeshScene mesh, OptRec opt, int iter, int n, const float* isects,
                               float* segments, cudaStream_t s);
cudaError_t configure_kernels();
// true when launch_bounce runs the persistent kernel for these parameters (everything except the
// linear scan over more than one shared-memory chunk of primitives)
bool bounce_is_persistent(const RenderParams& p);

// pt_gbvh.cu (host): nodes (4 float4 each, breadth first), the sphere / cube records in leaf order
// (pad0 = scene index) with their padded boxes, and the indices of the MESH primitives.  Empty nodes_out: nothing to index.
void build_geom_bvh(const std::vector<PrimRec>& prims, const double cam_pos[3], const double mesh_lo[3], const double mesh_hi[3],
                    bool have_mesh_bounds, std::vector<float4>& nodes_out, std::vector<PrimRec>& prims_out,
                    std::vector<float4>& boxes_out, std::vector<int>& mesh_prims_out, std::vector<uint4>& qnodes_out,
                    float qorg[3], float qcell[3]);

}  // namespace ptk
