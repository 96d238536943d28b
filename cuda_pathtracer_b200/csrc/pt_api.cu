// pt_api.cu -- the C ABI of include/b200_pt.h on top of the kernels in pt_kernels.cu.
//
// Host-side mirror of the reference's render driver (path_tracer/src/pathtrace.cu:113-177 init /
// free, :423-528 the per-iteration loop) without its per-bounce host work: no cudaMalloc / stream
// create / cudaDeviceSynchronize / D2H inside the loop; the live-path count of every bounce stays
// in device memory and `depth` bounce kernels are simply enqueued (tiles past the live range
// exit at once).  Several iterations are traced concurrently as independent batch slots of the
// same launches (each slot compacts in its own index space, so the per-(iteration, slot-index,
// depth) seeds -- and therefore the random paths -- are exactly the reference's).
#include <cstdlib>
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <new>
#include <utility>
#include <vector>

#include "../../include/b200_pt.h"
#include "../../include/b200_sc.h"
#include "pt_kernels.cuh"

using namespace ptk;

enum { B200_E_UNSUPPORTED = 10004 };   // B200_OK / BADARG / NODEVICE / NOMEM come from b200_sc.h

struct b200pt {
    int device = 0;
    int W = 0, H = 0, P = 0, depth = 0, tiles = 0, ptiles = 0;
    int n_prims = 0, n_mats = 0;
    CamRec cam{};
    b200pt_options opt{};
    int batch_cap = 0;                 // allocated batch slots
    PrimRec* d_prims = nullptr;
    MatRec* d_mats = nullptr;
    float4* d_pool = nullptr;          // 2 pools x 3 arrays x batch_cap x P
    int* d_counts = nullptr;           // (depth + 1) x batch_cap
    unsigned long long* d_desc = nullptr;
    int* d_tickets = nullptr;          // depth work-item counters of the persistent bounce kernel
    unsigned long long* d_segments = nullptr;
    float* d_image = nullptr;
    // split pipeline (material sort): hit records, radix-sort permutation ping-pong, digit histograms
    b200_sc_ctx* sc = nullptr;         // the scan library's workspace (exclusive scan of the histograms)
    float4* d_hit_n = nullptr;
    int* d_hit_k = nullptr;
    int* d_perm[2] = {nullptr, nullptr};
    int* d_hist = nullptr;
    int* d_hist_scanned = nullptr;
    int sort_batch_cap = 0;
    // first-bounce cache (valid while camera, geometry and the SSAA / DoF toggles stay put)
    float4* d_cache_n = nullptr;
    int* d_cache_k = nullptr;
    float2* d_hit_uv = nullptr;
    float2* d_cache_uv = nullptr;
    int* d_cache_perm = nullptr;       // bounce-0 order by material of the cached hits (material sort + cache)
    bool cache_valid = false;
    // triangle meshes / BVH / textures (b200pt_create_ext)
    BvhRec* d_nodes = nullptr;
    TriVert* d_tv = nullptr;
    TriAttr* d_ta = nullptr;
    TexRec* d_tex = nullptr;
    unsigned char* d_texels = nullptr;
    MeshScene mesh{};
    // geometry BVH over the sphere / cube primitives (pt_gbvh.cu); rebuilt whenever the primitives change
    float4* d_gnodes = nullptr;
    uint4* d_qnodes = nullptr;
    size_t qnodes_cap = 0;
    std::vector<uint4> host_qnodes;
    float4* d_gboxes = nullptr;
    PrimRec* d_gprims = nullptr;
    int* d_gmesh = nullptr;
    size_t gnodes_cap = 0, gboxes_cap = 0, gprims_cap = 0, gmesh_cap = 0;
    GeomBvh gb{};
    int n_indexed = 0;                 // sphere / cube primitives in the tree
    std::vector<float4> host_gnodes, host_gboxes;
    std::vector<PrimRec> host_gprims;
    std::vector<int> host_gmesh;
    unsigned epoch = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    uint64_t iterations = 0, launches = 0;
    float last_ms = 0.f;
    int last_batch = 0;
    std::vector<PrimRec> host_prims;   // staging for b200pt_update_scene
    // optional per-kernel timing of k_bounce (b200pt_set_profiling): one event pair brackets the
    // `depth` back-to-back bounce launches of a batch
    bool profiling = false;
    std::vector<cudaEvent_t> prof_events;
    size_t prof_used = 0;              // events in use (2 per batch)
    double bounce_ms = 0.0;
    uint64_t bounce_launches = 0;
    LaunchCfg cfg;                     // SM count / resident grids of the persistent kernels on h->device
};

namespace {

#define CK(x)                                    \
    do {                                         \
        cudaError_t e__ = (x);                   \
        if (e__ != cudaSuccess) return (int)e__; \
    } while (0)

// Every entry point runs on the device the handle was created on, whatever device the calling thread has
// current (renderers on several GPUs, or driven from several host threads, may coexist in one process);
// the caller's current device is restored on return.
struct DeviceGuard {
    int prev = -1;
    bool switched = false;
    cudaError_t err = cudaSuccess;
    explicit DeviceGuard(int device)
    {
        err = cudaGetDevice(&prev);
        if (err == cudaSuccess && prev != device) {
            err = cudaSetDevice(device);
            switched = err == cudaSuccess;
        }
    }
    ~DeviceGuard() { if (switched) cudaSetDevice(prev); }
};
#define ON_DEVICE(h)                                       \
    DeviceGuard guard__((h)->device);                       \
    if (guard__.err != cudaSuccess) return (int)guard__.err

int dev_alloc(void** p, size_t bytes)
{
    cudaError_t e = cudaMalloc(p, bytes);
    if (e != cudaSuccess) { cudaGetLastError(); return e == cudaErrorMemoryAllocation ? B200_E_NOMEM : (int)e; }
    return 0;
}

// struct Geom -> PrimRec: keep rows 0..2 of the three column-major matrices
void pack_prim(const b200pt_geom& g, PrimRec& r)
{
    for (int c = 0; c < 4; ++c)
        for (int row = 0; row < 3; ++row) {
            r.inv[c * 3 + row] = g.inverseTransform[c * 4 + row];
            r.fwd[c * 3 + row] = g.transform[c * 4 + row];
            r.itr[c * 3 + row] = g.invTranspose[c * 4 + row];
        }
    r.type = g.type;
    r.material = g.materialid;
    r.pad0 = g.type == B200PT_MESH ? g.tri_start_idx : 0;     // a MESH owns the triangle ids [pad0, pad1)
    r.pad1 = g.type == B200PT_MESH ? g.tri_end_idx : 0;
    if (g.type == B200PT_MESH)                                // its vertices are already in world space: `fwd` is free and
        for (int a = 0; a < 3; ++a) {                         // carries the world bounds meshIntersectionTest culls with
            r.fwd[a] = g.min_bound[a];
            r.fwd[3 + a] = g.max_bound[a];
        }
}

// Iterations in flight when the caller does not say: about 32 M paths (pools 3.2 GB + 1.2 GB of hit records where the
// split pipeline runs -- of 180 GB).  The late bounces of a wavefront are thin and their launches latency-bound, so
// the wider the wavefront the better, with diminishing returns: Cornell 800x800 14.3 / 15.3 / 15.45 / 15.4 G
// path-segments/s at 6 / 16 / 24 / 32 iterations, 1920x1080 glass 2582 / 3103 / 3192 spp/s at 2 / 8 / 16, config 5
// 154 / 162 / 166 spp/s at 1 / 2 / 4, mesh room 564 / 648 at 2 / 8 (profiles/r5_ab_variants.txt, r32-r33).
int default_batch(int P)
{
    int b = (int)((32u << 20) / (unsigned)(P > 0 ? P : 1));
    if (b < 1) b = 1;
    if (b > 32) b = 32;
    return b;
}

int ensure_batch(b200pt* h, int batch)
{
    if (batch <= h->batch_cap) return 0;
    cudaFree(h->d_pool); cudaFree(h->d_counts); cudaFree(h->d_desc);
    h->d_pool = nullptr; h->d_counts = nullptr; h->d_desc = nullptr; h->batch_cap = 0;
    int r = dev_alloc((void**)&h->d_pool, sizeof(float4) * 6 * (size_t)batch * h->P);
    if (r) return r;
    r = dev_alloc((void**)&h->d_counts, sizeof(int) * (size_t)(h->depth + 1) * batch);
    if (r) return r;
    r = dev_alloc((void**)&h->d_desc, sizeof(unsigned long long) * (size_t)batch * h->ptiles * PT_DESC_STRIDE);
    if (r) return r;
    CK(cudaMemset(h->d_counts, 0, sizeof(int) * (size_t)(h->depth + 1) * batch));
    CK(cudaMemset(h->d_desc, 0, sizeof(unsigned long long) * (size_t)batch * h->ptiles * PT_DESC_STRIDE));
    CK(cudaDeviceSynchronize());       // the memsets ran on the null stream: a non-blocking caller stream is not ordered behind them
    h->epoch = 0;
    h->batch_cap = batch;
    return 0;
}

int upload_meshes(b200pt* h, const b200pt_scene_ext* ext);

int grow_buffer(void** p, size_t* cap, size_t bytes)
{
    if (bytes <= *cap) return 0;
    cudaFree(*p);
    *p = nullptr; *cap = 0;
    int r = dev_alloc(p, bytes);
    if (!r) *cap = bytes;
    return r;
}

// (re)build the geometry BVH of the packed primitives and upload it on stream s
int upload_geom_bvh(b200pt* h, const b200pt_scene* sc, const std::vector<PrimRec>& prims, cudaStream_t s)
{
    const double cam[3] = {sc->camera.position[0], sc->camera.position[1], sc->camera.position[2]};
    double mlo[3] = {0, 0, 0}, mhi[3] = {0, 0, 0};
    bool have = false;
    for (int i = 0; i < sc->n_geoms; ++i) {
        const b200pt_geom& g = sc->geoms[i];
        if (g.type != B200PT_MESH || g.tri_end_idx <= g.tri_start_idx) continue;
        for (int a = 0; a < 3; ++a) {
            // (max_bound starts at FLT_MIN in the reference's loader, scene.cpp:115: may exceed the true maximum)
            const double lo = std::min((double)g.min_bound[a], (double)g.max_bound[a]), hi = std::max((double)g.min_bound[a], (double)g.max_bound[a]);
            mlo[a] = have ? std::min(mlo[a], lo) : lo;
            mhi[a] = have ? std::max(mhi[a], hi) : hi;
        }
        have = true;
    }
    float qorg[3] = {0, 0, 0}, qcell[3] = {1, 1, 1};
    build_geom_bvh(prims, cam, mlo, mhi, have, h->host_gnodes, h->host_gprims, h->host_gboxes, h->host_gmesh, h->host_qnodes, qorg, qcell);
    h->gb = GeomBvh{};
    h->n_indexed = (int)h->host_gprims.size();
    if (h->host_gnodes.empty()) return 0;
    int r = grow_buffer((void**)&h->d_gnodes, &h->gnodes_cap, sizeof(float4) * h->host_gnodes.size());
    if (!r) r = grow_buffer((void**)&h->d_gprims, &h->gprims_cap, sizeof(PrimRec) * h->host_gprims.size());
    if (!r) r = grow_buffer((void**)&h->d_gboxes, &h->gboxes_cap, sizeof(float4) * h->host_gboxes.size());
    if (!r) r = grow_buffer((void**)&h->d_gmesh, &h->gmesh_cap, sizeof(int) * (h->host_gmesh.size() + 1));
    if (!r && !h->host_qnodes.empty()) r = grow_buffer((void**)&h->d_qnodes, &h->qnodes_cap, sizeof(uint4) * h->host_qnodes.size());
    if (r) return r;
    if (!h->host_qnodes.empty()) {
        CK(cudaMemcpyAsync(h->d_qnodes, h->host_qnodes.data(), sizeof(uint4) * h->host_qnodes.size(), cudaMemcpyHostToDevice, s));
        h->gb.qnodes = h->d_qnodes;
        for (int a = 0; a < 3; ++a) { h->gb.qorg[a] = qorg[a]; h->gb.qcell[a] = qcell[a]; }
    }
    CK(cudaMemcpyAsync(h->d_gnodes, h->host_gnodes.data(), sizeof(float4) * h->host_gnodes.size(), cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(h->d_gprims, h->host_gprims.data(), sizeof(PrimRec) * h->host_gprims.size(), cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(h->d_gboxes, h->host_gboxes.data(), sizeof(float4) * h->host_gboxes.size(), cudaMemcpyHostToDevice, s));
    if (!h->host_gmesh.empty())
        CK(cudaMemcpyAsync(h->d_gmesh, h->host_gmesh.data(), sizeof(int) * h->host_gmesh.size(), cudaMemcpyHostToDevice, s));
    h->gb.nodes = h->d_gnodes;
    h->gb.prims = h->d_gprims;
    h->gb.boxes = h->d_gboxes;
    h->gb.crecs = h->d_gboxes + 2 * h->host_gprims.size();
    h->gb.n_small = h->n_indexed <= PT_SMALL_MAX ? h->n_indexed : 0;
    h->gb.mesh_prims = h->d_gmesh;
    h->gb.n_nodes = (int)(h->host_gnodes.size() / 4);
    h->gb.n_smem_nodes = std::min(h->gb.n_nodes, PT_SMEM_GNODES);
    h->gb.n_mesh_prims = (int)h->host_gmesh.size();
    h->gb.cap_mesh_walk = 1;
    for (int i = 0; i < sc->n_geoms; ++i)
        if (sc->geoms[i].type != B200PT_MESH && sc->materials[sc->geoms[i].materialid].texture_id != -1) h->gb.cap_mesh_walk = 0;
    return 0;
}

int sort_bins(const b200pt* h) { return h->n_mats < 256 ? h->n_mats : 256; }
int sort_passes(const b200pt* h)
{
    int passes = 1;
    while (passes < 4 && (h->n_mats - 1) >> (8 * passes)) ++passes;
    return passes;
}

// buffers of the split (material-sorted) pipeline, allocated on first use
int ensure_sort(b200pt* h, int batch)
{
    if (batch <= h->sort_batch_cap) return 0;
    cudaFree(h->d_hit_n); cudaFree(h->d_hit_k); cudaFree(h->d_perm[0]); cudaFree(h->d_perm[1]);
    cudaFree(h->d_hist); cudaFree(h->d_hist_scanned);
    h->d_hit_n = nullptr; h->d_hit_k = nullptr; h->d_perm[0] = h->d_perm[1] = nullptr;
    h->d_hist = h->d_hist_scanned = nullptr; h->sort_batch_cap = 0;
    const size_t paths = (size_t)batch * h->P;
    const size_t hist = (size_t)batch * sort_bins(h) * h->tiles + 4;
    cudaFree(h->d_hit_uv);
    h->d_hit_uv = nullptr;
    int r = dev_alloc((void**)&h->d_hit_n, sizeof(float4) * paths);
    if (!r) r = dev_alloc((void**)&h->d_hit_k, sizeof(int) * paths);
    if (!r) r = dev_alloc((void**)&h->d_hit_uv, sizeof(float2) * paths);
    if (!r) r = dev_alloc((void**)&h->d_perm[0], sizeof(int) * paths);
    if (!r && sort_passes(h) > 1) r = dev_alloc((void**)&h->d_perm[1], sizeof(int) * paths);
    if (!r) r = dev_alloc((void**)&h->d_hist, sizeof(int) * hist);
    if (!r) r = dev_alloc((void**)&h->d_hist_scanned, sizeof(int) * hist);
    if (r) return r;
    if (!h->sc) { r = b200_sc_ctx_create(&h->sc); if (r) return r; }
    h->sort_batch_cap = batch;
    return 0;
}

int ensure_cache(b200pt* h)
{
    if (h->d_cache_n) return 0;
    int r = dev_alloc((void**)&h->d_cache_n, sizeof(float4) * (size_t)h->P);
    if (!r) r = dev_alloc((void**)&h->d_cache_k, sizeof(int) * (size_t)h->P);
    if (!r) r = dev_alloc((void**)&h->d_cache_uv, sizeof(float2) * (size_t)h->P);
    if (!r) r = dev_alloc((void**)&h->d_cache_perm, sizeof(int) * (size_t)h->P);
    h->cache_valid = false;
    return r;
}

// fold the recorded event pairs into bounce_ms (synchronises on the last one)
int collect_profile(b200pt* h)
{
    if (!h->prof_used) return 0;
    CK(cudaEventSynchronize(h->prof_events[h->prof_used - 1]));
    for (size_t i = 0; i + 1 < h->prof_used; i += 2) {
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, h->prof_events[i], h->prof_events[i + 1]));
        h->bounce_ms += ms;
        h->bounce_launches += (uint64_t)h->depth;
    }
    h->prof_used = 0;
    return 0;
}

int prof_event(b200pt* h, cudaEvent_t* e)
{
    if (h->prof_used == h->prof_events.size()) {
        cudaEvent_t ne;
        CK(cudaEventCreate(&ne));
        h->prof_events.push_back(ne);
    }
    *e = h->prof_events[h->prof_used++];
    return 0;
}

void pack_camera(const b200pt_camera& c, CamRec& cam)
{
    cam.W = c.resolution[0];
    cam.H = c.resolution[1];
    for (int k = 0; k < 3; ++k) {
        cam.pos[k] = c.position[k];
        cam.view[k] = c.view[k];
        cam.up[k] = c.up[k];
        cam.right[k] = c.right[k];
    }
    cam.plx = c.pixelLength[0];
    cam.ply = c.pixelLength[1];
}

// geoms / materials / camera against the triangle, BVH-node and texture counts of the scene
int validate_scene_counts(const b200pt_scene* sc, int n_tris, int n_nodes, int n_tex)
{
    if (!sc || !sc->geoms || !sc->materials || sc->n_geoms <= 0 || sc->n_materials <= 0) return B200_E_BADARG;
    if (sc->camera.resolution[0] <= 0 || sc->camera.resolution[1] <= 0 || sc->trace_depth <= 0 || sc->trace_depth > 511)
        return B200_E_BADARG;
    for (int i = 0; i < sc->n_geoms; ++i) {
        const int type = sc->geoms[i].type;
        if (type != B200PT_SPHERE && type != B200PT_CUBE && type != B200PT_MESH) return B200_E_UNSUPPORTED;
        if (type == B200PT_MESH) {
            const int b = sc->geoms[i].tri_start_idx, e = sc->geoms[i].tri_end_idx;    // a range of triangle ids
            if (b < 0 || e < b || e > n_tris) return B200_E_BADARG;
            if (e > b && n_nodes == 0) return B200_E_UNSUPPORTED;                        // meshes are traced through the BVH
        }
        if (sc->geoms[i].materialid < 0 || sc->geoms[i].materialid >= sc->n_materials) return B200_E_BADARG;
    }
    for (int i = 0; i < sc->n_materials; ++i)
        if (sc->materials[i].texture_id < -1 || sc->materials[i].texture_id >= n_tex) return B200_E_BADARG;
    return 0;
}

int validate_scene(const b200pt_scene* sc, const b200pt_scene_ext* ext)
{
    const int n_tris = ext ? ext->n_triangles : 0, n_nodes = ext ? ext->n_bvh_nodes : 0, n_tex = ext ? ext->n_textures : 0;
    if (n_tris < 0 || n_nodes < 0 || n_tex < 0) return B200_E_BADARG;
    if ((n_tris && !ext->triangles) || (n_nodes && !ext->bvh) || (n_tex && !ext->textures)) return B200_E_BADARG;
    for (int i = 0; i < n_nodes; ++i) {
        const b200pt_bvh_node& nd = ext->bvh[i];
        if (nd.sub_areas > 0) {
            if (nd.first_area_idx < 0 || nd.first_area_idx > n_tris || nd.sub_areas > n_tris - nd.first_area_idx) return B200_E_BADARG;   // (no int overflow)
        } else if (nd.axis < 0 || nd.axis > 2 || nd.rchild_idx <= i || nd.rchild_idx >= n_nodes || i + 1 >= n_nodes) {
            return B200_E_BADARG;
        }
    }
    for (int i = 0; i < n_tex; ++i)
        if (ext->textures[i].width <= 0 || ext->textures[i].height <= 0 || ext->textures[i].components <= 0 ||
            !ext->textures[i].pixels)
            return B200_E_BADARG;
    return validate_scene_counts(sc, n_tris, n_nodes, n_tex);
}

// class Triangle / BVHTree / Texture -> the device records (pt_math.cuh)
int upload_meshes(b200pt* h, const b200pt_scene_ext* ext)
{
    const int n_tris = ext->n_triangles, n_nodes = ext->n_bvh_nodes, n_tex = ext->n_textures;
    if (n_nodes) {
        std::vector<BvhRec> nodes((size_t)n_nodes);
        for (int i = 0; i < n_nodes; ++i) {
            const b200pt_bvh_node& s = ext->bvh[i];
            BvhRec& d = nodes[i];
            for (int k = 0; k < 3; ++k) { d.lo[k] = s.min_corner[k]; d.hi[k] = s.max_corner[k]; }
            d.sub_areas = s.sub_areas;
            d.axis = s.sub_areas > 0 ? 0 : s.axis;
            d.first = s.sub_areas > 0 ? s.first_area_idx : 0;
            d.rchild = s.sub_areas > 0 ? 0 : s.rchild_idx;
            d.pad0 = d.pad1 = 0;
        }
        int r = dev_alloc((void**)&h->d_nodes, sizeof(BvhRec) * nodes.size());
        if (r) return r;
        CK(cudaMemcpy(h->d_nodes, nodes.data(), sizeof(BvhRec) * nodes.size(), cudaMemcpyHostToDevice));
    }
    if (n_tris) {
        std::vector<TriVert> tv((size_t)n_tris);
        std::vector<TriAttr> ta((size_t)n_tris);
        for (int i = 0; i < n_tris; ++i) {
            const b200pt_triangle& s = ext->triangles[i];
            for (int k = 0; k < 3; ++k) { tv[i].v0[k] = s.vertices[0][k]; tv[i].v1[k] = s.vertices[1][k]; tv[i].v2[k] = s.vertices[2][k]; }
            tv[i].id = s.id;
            tv[i].pad0 = tv[i].pad1 = 0;
            std::memcpy(ta[i].uv, s.uv, sizeof(s.uv));
            std::memcpy(ta[i].n, s.normals, sizeof(s.normals));
            ta[i].pad = 0.f;
        }
        int r = dev_alloc((void**)&h->d_tv, sizeof(TriVert) * tv.size());
        if (!r) r = dev_alloc((void**)&h->d_ta, sizeof(TriAttr) * ta.size());
        if (r) return r;
        CK(cudaMemcpy(h->d_tv, tv.data(), sizeof(TriVert) * tv.size(), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(h->d_ta, ta.data(), sizeof(TriAttr) * ta.size(), cudaMemcpyHostToDevice));
    }
    if (n_tex) {
        std::vector<TexRec> recs((size_t)n_tex);
        std::vector<unsigned char> pool;
        for (int i = 0; i < n_tex; ++i) {
            const b200pt_texture& t = ext->textures[i];
            const size_t bytes = (size_t)t.width * t.height * t.components;
            recs[i].w = t.width; recs[i].h = t.height; recs[i].comp = t.components;
            if (pool.size() + bytes > 0xffffffffull) return B200_E_BADARG;     // TexRec::offset is 32 bits
            recs[i].offset = (unsigned)pool.size();
            pool.insert(pool.end(), t.pixels, t.pixels + bytes);
            while (pool.size() % 16) pool.push_back(0);
        }
        int r = dev_alloc((void**)&h->d_tex, sizeof(TexRec) * recs.size());
        if (!r) r = dev_alloc((void**)&h->d_texels, pool.size());
        if (r) return r;
        CK(cudaMemcpy(h->d_tex, recs.data(), sizeof(TexRec) * recs.size(), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(h->d_texels, pool.data(), pool.size(), cudaMemcpyHostToDevice));
    }
    h->mesh.nodes = h->d_nodes;
    h->mesh.tv = h->d_tv;
    h->mesh.ta = h->d_ta;
    h->mesh.tex = h->d_tex;
    h->mesh.texels = h->d_texels;
    h->mesh.n_nodes = n_tris ? n_nodes : 0;
    h->mesh.n_tris = n_tris;
    h->mesh.n_tex = n_tex;
    h->mesh.n_smem_nodes = h->mesh.n_nodes < PT_SMEM_NODES ? h->mesh.n_nodes : PT_SMEM_NODES;
    // depth of the tree (levels): below BVH_MAX_SEARCH_DEPTH the reference's 64-entry stack can never
    // fill up, so its overflow behaviour is unreachable and the deferred / pruned walk is exact
    int levels = 0;
    if (h->mesh.n_nodes) {
        std::vector<std::pair<int, int>> todo{{0, 1}};
        while (!todo.empty()) {
            const std::pair<int, int> at = todo.back();
            todo.pop_back();
            levels = std::max(levels, at.second);
            if (levels > 4 * BVH_MAX_SEARCH_DEPTH) break;
            const b200pt_bvh_node& nd = ext->bvh[at.first];
            if (nd.sub_areas > 0) continue;
            todo.push_back({at.first + 1, at.second + 1});
            todo.push_back({nd.rchild_idx, at.second + 1});
        }
    }
    h->mesh.literal_walk = levels >= BVH_MAX_SEARCH_DEPTH ? 1 : 0;
    return 0;
}

}  // namespace

extern "C" {

void b200pt_default_options(b200pt_options* o)
{
    if (!o) return;
    o->russian_roulette = 1;      // utilities.h:23-33
    o->sort_by_material = 0;
    o->ssaa = 1;
    o->dof = 1;
    o->aperture = 0.1f;
    o->focal_len = 10.0f;
    o->first_bounce_cache = 0;
    o->batch = 0;
    o->geom_bvh = 0;
    o->mesh_brute_force = 0;
    o->mesh_no_bbox = 0;
    o->use_thrust_partition = 0;
}

int b200pt_create(const b200pt_scene* sc, b200pt** out) { return b200pt_create_ext(sc, nullptr, out); }

int b200pt_create_ext(const b200pt_scene* sc, const b200pt_scene_ext* ext, b200pt** out)
{
    if (!out) return B200_E_BADARG;
    const int bad = validate_scene(sc, ext);
    if (bad) return bad;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) { cudaGetLastError(); return B200_E_NODEVICE; }

    b200pt* h = new (std::nothrow) b200pt();
    if (!h) return B200_E_NOMEM;
    // (from here on every failure goes through b200pt_destroy: nothing allocated so far is leaked)
    int r = (int)cudaGetDevice(&h->device);
    if (!r) r = (int)configure_kernels();
    if (r) { b200pt_destroy(h); return r; }
    h->W = sc->camera.resolution[0];
    h->H = sc->camera.resolution[1];
    h->P = h->W * h->H;
    h->depth = sc->trace_depth;
    h->tiles = (h->P + PT_THREADS - 1) / PT_THREADS;
    h->ptiles = (h->P + PT_PTILE - 1) / PT_PTILE;
    h->n_prims = sc->n_geoms;
    h->n_mats = sc->n_materials;
    pack_camera(sc->camera, h->cam);
    b200pt_default_options(&h->opt);

    std::vector<PrimRec> prims(sc->n_geoms);
    for (int i = 0; i < sc->n_geoms; ++i) pack_prim(sc->geoms[i], prims[i]);
    r = dev_alloc((void**)&h->d_prims, sizeof(PrimRec) * prims.size());
    if (!r) r = dev_alloc((void**)&h->d_mats, sizeof(MatRec) * (size_t)sc->n_materials);
    if (!r) r = dev_alloc((void**)&h->d_image, sizeof(float) * 3 * (size_t)h->P);
    if (!r) r = dev_alloc((void**)&h->d_segments, sizeof(unsigned long long));
    if (!r) r = dev_alloc((void**)&h->d_tickets, sizeof(int) * 2 * (size_t)(h->depth > 0 ? h->depth : 1));
    static_assert(sizeof(b200pt_material) == sizeof(MatRec), "material layouts must agree");
    if (!r) r = (int)cudaMemcpy(h->d_prims, prims.data(), sizeof(PrimRec) * prims.size(), cudaMemcpyHostToDevice);
    if (!r) r = (int)cudaMemcpy(h->d_mats, sc->materials, sizeof(MatRec) * (size_t)sc->n_materials, cudaMemcpyHostToDevice);
    if (!r) r = (int)cudaEventCreate(&h->ev0);
    if (!r) r = (int)cudaEventCreate(&h->ev1);
    if (!r) r = b200pt_reset(h);
    if (!r && ext) r = upload_meshes(h, ext);
    if (!r) r = upload_geom_bvh(h, sc, prims, nullptr);
    if (!r && cudaStreamSynchronize(nullptr) != cudaSuccess) r = (int)cudaGetLastError();
    if (r) { b200pt_destroy(h); return r; }
    *out = h;
    return 0;
}

int b200pt_destroy(b200pt* h)
{
    if (!h) return 0;
    DeviceGuard guard(h->device);
    cudaFree(h->d_prims); cudaFree(h->d_mats); cudaFree(h->d_pool); cudaFree(h->d_counts);
    cudaFree(h->d_desc); cudaFree(h->d_tickets); cudaFree(h->d_segments); cudaFree(h->d_image);
    cudaFree(h->d_hit_n); cudaFree(h->d_hit_k); cudaFree(h->d_perm[0]); cudaFree(h->d_perm[1]);
    cudaFree(h->d_hist); cudaFree(h->d_hist_scanned); cudaFree(h->d_cache_n); cudaFree(h->d_cache_k);
    cudaFree(h->d_hit_uv); cudaFree(h->d_cache_uv); cudaFree(h->d_cache_perm);
    cudaFree(h->d_nodes); cudaFree(h->d_tv); cudaFree(h->d_ta); cudaFree(h->d_tex); cudaFree(h->d_texels);
    cudaFree(h->d_gnodes); cudaFree(h->d_qnodes); cudaFree(h->d_gboxes); cudaFree(h->d_gprims); cudaFree(h->d_gmesh);
    b200_sc_ctx_destroy(h->sc);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    for (cudaEvent_t e : h->prof_events) cudaEventDestroy(e);
    delete h;
    return 0;
}

int b200pt_set_options(b200pt* h, const b200pt_options* o)
{
    if (!h || !o || o->batch < 0 || o->batch > PT_MAX_BATCH || o->geom_bvh < -1 || o->geom_bvh > 1) return B200_E_BADARG;
    // the cached first-bounce hits depend on the camera rays: any change of what shapes them drops the cache
    if (o->ssaa != h->opt.ssaa || o->dof != h->opt.dof || o->first_bounce_cache != h->opt.first_bounce_cache ||
        o->mesh_brute_force != h->opt.mesh_brute_force || o->mesh_no_bbox != h->opt.mesh_no_bbox ||
        o->sort_by_material != h->opt.sort_by_material)      // (the cache also keeps the hits' order by material)
        h->cache_valid = false;
    h->opt = *o;
    return 0;
}

int b200pt_reset(b200pt* h)
{
    if (!h) return B200_E_BADARG;
    ON_DEVICE(h);
    CK(cudaMemset(h->d_image, 0, sizeof(float) * 3 * (size_t)h->P));
    CK(cudaMemset(h->d_segments, 0, sizeof(unsigned long long)));
    CK(cudaDeviceSynchronize());       // (null-stream memsets: make them visible to launches on non-blocking streams)
    h->iterations = 0;
    h->launches = 0;
    h->prof_used = 0;
    h->bounce_ms = 0.0;
    h->bounce_launches = 0;
    return 0;
}

// Re-upload geometry, materials and camera (same counts, resolution and depth): what the
// reference does through pathtraceFree + pathtraceInit whenever the camera or a visual setting
// changes (main.cpp:142-146), without re-allocating.  The accumulation image is NOT cleared.
int b200pt_update_scene(b200pt* h, const b200pt_scene* sc, void* stream)
{
    if (!h) return B200_E_BADARG;
    ON_DEVICE(h);
    {   // triangles, BVH and textures stay as uploaded: validate against their counts
        const int bad = validate_scene_counts(sc, h->mesh.n_tris, h->mesh.n_nodes, h->mesh.n_tex);
        if (bad) return bad;
    }
    if (sc->n_geoms != h->n_prims || sc->n_materials != h->n_mats || sc->camera.resolution[0] != h->W ||
        sc->camera.resolution[1] != h->H || sc->trace_depth != h->depth)
        return B200_E_BADARG;
    h->host_prims.resize((size_t)sc->n_geoms);
    for (int i = 0; i < sc->n_geoms; ++i) pack_prim(sc->geoms[i], h->host_prims[i]);
    cudaStream_t s = (cudaStream_t)stream;
    CK(cudaMemcpyAsync(h->d_prims, h->host_prims.data(), sizeof(PrimRec) * h->host_prims.size(), cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(h->d_mats, sc->materials, sizeof(MatRec) * (size_t)sc->n_materials, cudaMemcpyHostToDevice, s));
    {
        const int r = upload_geom_bvh(h, sc, h->host_prims, s);
        if (r) return r;
    }
    CK(cudaStreamSynchronize(s));      // host_prims / caller memory may be reused on return
    pack_camera(sc->camera, h->cam);
    h->cache_valid = false;
    return 0;
}

int b200pt_set_profiling(b200pt* h, int enable)
{
    if (!h) return B200_E_BADARG;
    ON_DEVICE(h);
    int r = collect_profile(h);
    if (r) return r;
    h->profiling = enable != 0;
    return 0;
}

int b200pt_render(b200pt* h, int first_iter, int n_iters, int iter_stride, void* stream, int sync)
{
    if (!h || n_iters < 0 || iter_stride < 1 || first_iter < 0) return B200_E_BADARG;
    ON_DEVICE(h);
    cudaStream_t s = (cudaStream_t)stream;
    // meshIntersectionTest instead of the BVH walk: the reference's own loop over all primitives (the geometry
    // BVH handles MESH primitives through the one shared walk, which this mode does not have)
    const bool brute = h->opt.mesh_brute_force != 0 && h->mesh.n_tris > 0;
    const bool use_gb = !brute && h->gb.n_nodes > 0 && (h->opt.geom_bvh > 0 || (h->opt.geom_bvh == 0 && h->n_indexed >= PT_GBVH_AUTO));
    const bool sort_on = h->opt.sort_by_material != 0;
    // Scenes whose closest hit is a tree walk (triangle meshes; the geometry BVH proper of a many-primitive
    // scene) run closest hit and shading as TWO kernels, the hit records going through HBM (36 B per path) --
    // the material-sort pipeline without the sort.  The walk alone needs fewer registers than the fused
    // kernel and its long, uneven loops no longer hold the shading state hostage: config 5 89.5 -> 102.3
    // spp/s, mesh room 355 -> 411 (profiles/r5_ab_variants.txt).  Same hits, same slot order, same seeds: the
    // image does not change.  Small mesh-free scenes (Cornell) keep the fused persistent kernel.
#ifndef PT_FUSE_RAYGEN
#define PT_FUSE_RAYGEN 1     // bounce 0 of the persistent fused kernel generates the camera rays itself (no k_raygen launch)
#endif
#ifndef PT_SPLIT_MODE
#define PT_SPLIT_MODE 2      // 0 = never, 1 = scenes with meshes, 2 = + geometry BVH proper, 3 = always (experiment)
#endif
    const bool split_only = !sort_on && ((PT_SPLIT_MODE >= 1 && h->mesh.n_nodes > 0) ||
                                         (PT_SPLIT_MODE >= 2 && use_gb && h->gb.n_small == 0) || PT_SPLIT_MODE >= 3);
    const bool cache_on = h->opt.first_bounce_cache && !h->opt.ssaa && !h->opt.dof;

    // iterations in flight: never more than this call traces (a one-iteration call, e.g. the reference-shaped
    // pathtrace(), then allocates one slot, not 32), and halved until the pools fit into device memory
    int batch = h->opt.batch > 0 ? h->opt.batch : default_batch(h->P);
    if (n_iters > 0 && batch > n_iters) batch = n_iters;
    int r = 0;
    for (;;) {
        r = ensure_batch(h, batch);
        if (!r && (sort_on || split_only)) r = ensure_sort(h, batch);
        if (r != B200_E_NOMEM || batch == 1) break;
        batch = (batch + 1) / 2;
    }
    if (r) return r;
    if (cache_on) { r = ensure_cache(h); if (r) return r; }

    RenderParams p{};
    const size_t plane = (size_t)h->batch_cap * h->P;
    Pool pools[2];
    for (int k = 0; k < 2; ++k) {
        pools[k].a = h->d_pool + (size_t)(3 * k + 0) * plane;
        pools[k].b = h->d_pool + (size_t)(3 * k + 1) * plane;
        pools[k].c = h->d_pool + (size_t)(3 * k + 2) * plane;
    }
    p.counts = h->d_counts;
    p.desc = h->d_desc;
    p.segments = h->d_segments;
    p.image = h->d_image;
    p.prims = h->d_prims;
    p.mats = h->d_mats;
    p.n_prims = h->n_prims;
    p.n_mats = h->n_mats;
    if (use_gb) p.gb = h->gb;
    p.cam = h->cam;
    p.opt.rr = h->opt.russian_roulette;
    p.opt.ssaa = h->opt.ssaa;
    p.opt.dof = h->opt.dof;
    p.opt.focal = h->opt.focal_len;
    p.opt.aperture = h->opt.aperture;
    p.depth = h->depth;
    p.P = h->P;
    p.tiles = h->tiles;
    p.ptiles = h->ptiles;
    p.tickets = h->d_tickets;
    p.stride = iter_stride;

    p.hit_n = h->d_hit_n;
    p.hit_k = h->d_hit_k;
    p.hist = h->d_hist;
    p.hist_scanned = h->d_hist_scanned;
    p.cache_n = h->d_cache_n;
    p.cache_k = h->d_cache_k;
    p.hit_uv = h->d_hit_uv;
    p.cache_uv = h->d_cache_uv;
    p.mesh = h->mesh;
    p.mesh.brute = brute ? (h->opt.mesh_no_bbox ? 2 : 1) : 0;

    CK(cudaEventRecord(h->ev0, s));
    for (int done = 0; done < n_iters; done += batch) {
        const int nb = n_iters - done < batch ? n_iters - done : batch;
        p.batch = nb;
        p.iter0 = first_iter + done * iter_stride;
        p.in = pools[0];
        p.out = pools[1];
        // the plain fused pipeline on the persistent kernel generates the camera rays inside bounce 0 (RenderParams::raygen)
        p.cache_mode = 0;
        const bool fuse_raygen = PT_FUSE_RAYGEN && !split_only && !sort_on && !cache_on && bounce_is_persistent(p);
        if (fuse_raygen) CK(launch_batch_init(p, s));
        else CK(launch_raygen(p, s));
        if (h->profiling) {
            if (h->prof_used >= 8192) { r = collect_profile(h); if (r) return r; }
            cudaEvent_t e;
            r = prof_event(h, &e);
            if (r) return r;
            CK(cudaEventRecord(e, s));
        }
        for (int b = 0; b < h->depth; ++b) {
            p.bounce = b;
            if (++h->epoch >= (1u << 30)) {
                CK(cudaMemsetAsync(h->d_desc, 0, sizeof(unsigned long long) * (size_t)h->batch_cap * h->ptiles * PT_DESC_STRIDE, s));
                h->epoch = 1;
            }
            p.epoch = h->epoch;
            p.cache_mode = (b == 0 && cache_on) ? (h->cache_valid ? 2 : 1) : 0;
            p.raygen = (b == 0 && fuse_raygen) ? 1 : 0;
            p.perm_shared = 0;
            p.sort_bins = 0;
            if (split_only) {
                if (p.cache_mode != 2) { CK(launch_isect(p, h->cfg, s)); h->launches += 1; }
                p.perm_in = nullptr;
                CK(launch_shade(p, h->cfg, s));
                h->launches += 1;
            } else if (!sort_on) {
                CK(launch_bounce(p, h->cfg, s));
                h->launches += 1;
            } else {
                // computeIntersections -> sort_by_key(materialId) -> shadeMaterials (pathtrace.cu:470-497)
                const int bins = sort_bins(h), passes = sort_passes(h);
                bool hist_done = false;
#ifdef PT_NO_FUSED_HIST
                p.sort_bins = 0;
#else
                p.sort_bins = passes == 1 ? bins : 0;        // one radix pass: the closest-hit kernel does the counting half where it can
#endif
                if (p.cache_mode != 2) { CK(launch_isect(p, h->cfg, s, &hist_done)); h->launches += 1; }
                if (p.cache_mode == 2) {
                    // The cached bounce-0 hits are the same every iteration, and so is their stable order by material:
                    // it was stored when the cache was filled, one permutation for all slots.  No sort at all here.
                    p.perm_in = h->d_cache_perm;
                    p.perm_shared = 1;
                } else {
                    for (int pass = 0; pass < passes; ++pass) {
                        p.perm_in = pass == 0 ? nullptr : h->d_perm[(pass - 1) & 1];
                        p.perm_out = h->d_perm[pass & 1];
                        if (!(pass == 0 && hist_done)) { CK(launch_sort_hist(p, h->cfg, pass, bins, s)); h->launches += 1; }
                        r = b200_scan_i32_dev(h->sc, nb * bins * h->tiles, h->d_hist_scanned, h->d_hist, B200_SCAN_LOOKBACK,
                                              nullptr, nullptr, s);
                        if (r) return r;
                        CK(launch_sort_scatter(p, h->cfg, pass, bins, pass == 0 && hist_done, s));
                        h->launches += 1 + (uint64_t)b200_sc_last_launches(h->sc);
                    }
                    p.perm_in = h->d_perm[(passes - 1) & 1];
                    if (p.cache_mode == 1)                   // slot 0's order of the cached hits, kept beside them
                        CK(cudaMemcpyAsync(h->d_cache_perm, p.perm_in, sizeof(int) * (size_t)h->P, cudaMemcpyDeviceToDevice, s));
                }
                CK(launch_shade(p, h->cfg, s));
                h->launches += 1;
            }
            if (p.cache_mode == 1) h->cache_valid = true;   // later launches are stream-ordered behind the fill
            const Pool t = p.in; p.in = p.out; p.out = t;
        }
        if (h->profiling) {
            cudaEvent_t e;
            r = prof_event(h, &e);
            if (r) return r;
            CK(cudaEventRecord(e, s));
        }
        h->launches += 1;
        h->last_batch = nb;
    }
    CK(cudaEventRecord(h->ev1, s));
    h->iterations += (uint64_t)n_iters;
    if (sync) {
        CK(cudaEventSynchronize(h->ev1));
        CK(cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1));
    }
    return 0;
}

int b200pt_tonemap_u8(b200pt* h, int iter_divisor, void* dev_rgba8, void* stream)
{
    if (!h || !dev_rgba8 || iter_divisor <= 0) return B200_E_BADARG;
    ON_DEVICE(h);
    CK(launch_tonemap(h->d_image, h->P, iter_divisor, (uchar4*)dev_rgba8, (cudaStream_t)stream));
    h->launches += 1;
    return 0;
}

int b200pt_tonemap_buffer_u8(b200pt* h, const void* dev_sum_rgb, int iter_divisor, void* dev_rgba8, void* stream)
{
    if (!h || !dev_sum_rgb || !dev_rgba8 || iter_divisor <= 0) return B200_E_BADARG;
    ON_DEVICE(h);
    CK(launch_tonemap((const float*)dev_sum_rgb, h->P, iter_divisor, (uchar4*)dev_rgba8, (cudaStream_t)stream));
    h->launches += 1;
    return 0;
}

int b200pt_read_sum(b200pt* h, float* host_rgb)
{
    if (!h || !host_rgb) return B200_E_BADARG;
    ON_DEVICE(h);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(host_rgb, h->d_image, sizeof(float) * 3 * (size_t)h->P, cudaMemcpyDeviceToHost));
    return 0;
}

int b200pt_write_sum(b200pt* h, const float* host_rgb)
{
    if (!h || !host_rgb) return B200_E_BADARG;
    ON_DEVICE(h);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(h->d_image, host_rgb, sizeof(float) * 3 * (size_t)h->P, cudaMemcpyHostToDevice));
    CK(cudaDeviceSynchronize());
    return 0;
}

int b200pt_sum_device_ptr(b200pt* h, void** dev_ptr, size_t* n_floats)
{
    if (!h || !dev_ptr) return B200_E_BADARG;
    *dev_ptr = h->d_image;
    if (n_floats) *n_floats = 3 * (size_t)h->P;
    return 0;
}

int b200pt_get_stats(b200pt* h, b200pt_stats* st)
{
    if (!h || !st) return B200_E_BADARG;
    ON_DEVICE(h);
    CK(cudaDeviceSynchronize());
    unsigned long long seg = 0;
    CK(cudaMemcpy(&seg, h->d_segments, sizeof(seg), cudaMemcpyDeviceToHost));
    st->path_segments = seg;
    st->iterations = h->iterations;
    st->kernel_launches = h->launches;
    if (h->ev1 && cudaEventQuery(h->ev1) == cudaSuccess && h->iterations) cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
    cudaGetLastError();
    st->last_render_ms = h->last_ms;
    { int r = collect_profile(h); if (r) return r; }
    st->bounce_kernel_ms = h->bounce_ms;
    st->bounce_kernel_launches = h->bounce_launches;
    st->traced_depth = 0;
    if (h->batch_cap && h->last_batch) {
        // TracedDepth (pathtrace.cu:510): bounces the last iteration ran before its pool emptied
        std::vector<int> counts((size_t)(h->depth + 1) * h->last_batch);
        CK(cudaMemcpy(counts.data(), h->d_counts, sizeof(int) * counts.size(), cudaMemcpyDeviceToHost));
        int d = 0;
        for (int b = 0; b < h->depth; ++b)
            if (counts[(size_t)b * h->last_batch + (h->last_batch - 1)] > 0) d = b + 1;
        st->traced_depth = d;
    }
    return 0;
}

static void fill_params(const b200pt* h, RenderParams& p)
{
    p.cam = h->cam;
    p.opt.rr = h->opt.russian_roulette;
    p.opt.ssaa = h->opt.ssaa;
    p.opt.dof = h->opt.dof;
    p.opt.focal = h->opt.focal_len;
    p.opt.aperture = h->opt.aperture;
    p.depth = h->depth;
    p.P = h->P;
    p.tiles = h->tiles;
}

int b200pt_debug_raygen(b200pt* h, int iter, void* d_segments, void* stream)
{
    if (!h || !d_segments) return B200_E_BADARG;
    ON_DEVICE(h);
    RenderParams p{};
    fill_params(h, p);
    CK(launch_debug_raygen(p, iter, (float*)d_segments, (cudaStream_t)stream));
    h->launches += 1;
    return 0;
}

int b200pt_debug_intersect(b200pt* h, int n, const void* d_segments, void* d_isects, void* stream)
{
    if (!h || n < 0 || (n > 0 && (!d_segments || !d_isects))) return B200_E_BADARG;
    if (n == 0) return 0;
    ON_DEVICE(h);
    MeshScene mesh = h->mesh;
    mesh.brute = h->opt.mesh_brute_force != 0 && mesh.n_tris > 0 ? (h->opt.mesh_no_bbox ? 2 : 1) : 0;
    CK(launch_debug_intersect(h->d_prims, h->n_prims, mesh, n, (const float*)d_segments, (float*)d_isects, (cudaStream_t)stream));
    h->launches += 1;
    return 0;
}

int b200pt_debug_shade(b200pt* h, int iter, int n, const void* d_isects, void* d_segments, void* stream)
{
    if (!h || n < 0 || (n > 0 && (!d_segments || !d_isects))) return B200_E_BADARG;
    if (n == 0) return 0;
    ON_DEVICE(h);
    RenderParams p{};
    fill_params(h, p);
    CK(launch_debug_shade(h->d_mats, h->mesh, p.opt, iter, n, (const float*)d_isects, (float*)d_segments, (cudaStream_t)stream));
    h->launches += 1;
    return 0;
}

const char* b200pt_error_string(int code)
{
    switch (code) {
    case B200_OK: return "ok";
    case B200_E_BADARG: return "b200pt: bad argument";
    case B200_E_NODEVICE: return "b200pt: no CUDA device (this library has no CPU fallback)";
    case B200_E_NOMEM: return "b200pt: out of device memory";
    case B200_E_UNSUPPORTED: return "b200pt: scene feature not supported (triangle meshes, textures)";
    default: return cudaGetErrorString((cudaError_t)code);
    }
}

}  // extern "C"
