// TEST / BASELINE INFRASTRUCTURE -- not product code.
//
// Launches the reference's OWN kernels (generateRayFromCamera, computeIntersections,
// shadeMaterials: path_tracer/src/pathtrace.cu:184-344, compiled unmodified from /root/reference
// with -rdc=true, so the __global__ symbols are linkable) one stage at a time on arrays read from
// files, so that the per-stage parity tests (tests/test_pt_gpu.py) can compare this repo's
// kernels with them bit for bit on identical inputs (SURVEY.md 8c tier T1).
//
//   ref_stage_probe raygen    SCENE.bin ITER OUT_segments.bin            [flags]
//   ref_stage_probe intersect SCENE.bin IN_segments.bin OUT_isects.bin   [flags]
//   ref_stage_probe shade     SCENE.bin ITER IN_segments.bin IN_isects.bin OUT_segments.bin [flags]
//   flags: --no-ssaa --no-dof --no-rr --no-bvh --no-bbox --aperture A --focal F
// SCENE.bin is a "B2PTSCN1" dump (oracle/ref_scene_dump.cpp); segments are the reference's
// 48-byte PathSegment, isects its 28-byte ShadeableIntersection.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "scene.h"
#include "sceneStructs.h"
#include "utilities.h"

Scene::~Scene() {}

// pathtrace.cu:31-42 (the class is file-local there; same layout here)
class Settings {
public:
    bool enable_RussianRoulette;
    bool enable_SortbyMaterial;
    bool use_Thrust;
    bool use_BoundingBoxTest;
    bool use_BVHtree;
    bool enable_SSAA;
    bool enable_DoF;
    float focal_dist;
    float aperture;
};

__global__ void generateRayFromCamera(Camera cam, int iter, int traceDepth, PathSegment* pathSegments, Settings* settings);
__global__ void computeIntersections(int depth, int num_paths, PathSegment* pathSegments, Geom* geoms, int geoms_size,
                                     ShadeableIntersection* intersections, Triangle* triangles, BoundingBox* bounds,
                                     BVHTree* bvh_tree, Settings* settings);
__global__ void shadeMaterials(int iter, int num_paths, Settings* settings, ShadeableIntersection* shadeableIntersections,
                               PathSegment* pathSegments, Material* materials, Texture* textures);

template <class T>
static std::vector<T> read_file(const char* path)
{
    FILE* f = fopen(path, "rb");
    if (!f) { perror(path); exit(1); }
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    std::vector<T> v(n / sizeof(T));
    if (fread(v.data(), sizeof(T), v.size(), f) != v.size()) { fprintf(stderr, "short read %s\n", path); exit(1); }
    fclose(f);
    return v;
}

template <class T>
static void write_file(const char* path, const T* d_ptr, size_t n)
{
    std::vector<T> h(n);
    cudaMemcpy(h.data(), d_ptr, n * sizeof(T), cudaMemcpyDeviceToHost);
    FILE* f = fopen(path, "wb");
    if (!f) { perror(path); exit(1); }
    fwrite(h.data(), sizeof(T), n, f);
    fclose(f);
}

template <class T>
static T* upload(const std::vector<T>& v)
{
    T* d = nullptr;
    cudaMalloc(&d, (v.size() ? v.size() : 1) * sizeof(T));
    cudaMemcpy(d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice);
    return d;
}

int main(int argc, char** argv)
{
    if (argc < 4) { fprintf(stderr, "usage: see the header of oracle/ref_stage_probe.cu\n"); return 1; }
    const std::string mode = argv[1];
    std::vector<char> raw = read_file<char>(argv[2]);
    const bool v2 = memcmp(raw.data(), "B2PTSCN2", 8) == 0;
    if (!v2 && memcmp(raw.data(), "B2PTSCN1", 8) != 0) { fprintf(stderr, "bad scene dump\n"); return 1; }
    int hdr[4];
    memcpy(hdr, raw.data() + 8, 16);
    Camera cam;
    memcpy(&cam, raw.data() + 24, sizeof(Camera));
    std::vector<Geom> geoms(hdr[0]);
    memcpy(geoms.data(), raw.data() + 24 + sizeof(Camera), sizeof(Geom) * hdr[0]);
    std::vector<Material> mats(hdr[1]);
    memcpy(mats.data(), raw.data() + 24 + sizeof(Camera) + sizeof(Geom) * hdr[0], sizeof(Material) * hdr[1]);
    const int depth = hdr[2];
    // triangle meshes / BVH / textures (B2PTSCN2), uploaded as pathtraceInit does (pathtrace.cu:137-166)
    Triangle* d_tris = nullptr;
    BVHTree* d_bvh = nullptr;
    Texture* d_textures = nullptr;
    if (v2) {
        const char* at = raw.data() + 24 + sizeof(Camera) + sizeof(Geom) * hdr[0] + sizeof(Material) * hdr[1];
        int ext[4];
        memcpy(ext, at, 16);
        at += 16;
        std::vector<Triangle> tris(ext[0]);
        memcpy((void*)tris.data(), at, sizeof(Triangle) * ext[0]);
        at += sizeof(Triangle) * ext[0];
        std::vector<BVHTree> nodes(ext[1]);
        memcpy((void*)nodes.data(), at, sizeof(BVHTree) * ext[1]);
        at += sizeof(BVHTree) * ext[1];
        std::vector<Texture> texs(ext[2]);
        for (int i = 0; i < ext[2]; ++i) {
            int th[4];
            memcpy(th, at, 16);
            at += 16;
            texs[i].id = i; texs[i].width = th[0]; texs[i].height = th[1]; texs[i].components = th[2];
            const size_t bytes = (size_t)th[0] * th[1] * th[2];
            cudaMalloc(&texs[i].dev_image, bytes);
            cudaMemcpy(texs[i].dev_image, at, bytes, cudaMemcpyHostToDevice);
            at += bytes;
        }
        if (ext[0]) { cudaMalloc(&d_tris, sizeof(Triangle) * ext[0]); cudaMemcpy(d_tris, tris.data(), sizeof(Triangle) * ext[0], cudaMemcpyHostToDevice); }
        if (ext[0] && ext[1]) { cudaMalloc(&d_bvh, sizeof(BVHTree) * ext[1]); cudaMemcpy(d_bvh, nodes.data(), sizeof(BVHTree) * ext[1], cudaMemcpyHostToDevice); }
        if (ext[2]) { cudaMalloc(&d_textures, sizeof(Texture) * ext[2]); cudaMemcpy(d_textures, texs.data(), sizeof(Texture) * ext[2], cudaMemcpyHostToDevice); }
    }

    Settings st{};
    st.enable_RussianRoulette = true; st.use_BoundingBoxTest = true; st.use_BVHtree = true;
    st.enable_SSAA = true; st.enable_DoF = true; st.aperture = 0.1f; st.focal_dist = 10.0f;
    const int first_flag = mode == "raygen" ? 5 : (mode == "intersect" ? 5 : 7);
    for (int i = first_flag; i < argc; ++i) {
        std::string a = argv[i];
        if (a == "--no-ssaa") st.enable_SSAA = false;
        else if (a == "--no-dof") st.enable_DoF = false;
        else if (a == "--no-rr") st.enable_RussianRoulette = false;
        else if (a == "--no-bvh") st.use_BVHtree = false;
        else if (a == "--no-bbox") st.use_BoundingBoxTest = false;
        else if (a == "--aperture" && i + 1 < argc) st.aperture = (float)atof(argv[++i]);
        else if (a == "--focal" && i + 1 < argc) st.focal_dist = (float)atof(argv[++i]);
        else { fprintf(stderr, "unknown flag %s\n", a.c_str()); return 1; }
    }
    Settings* d_st = nullptr;
    cudaMalloc(&d_st, sizeof(Settings));
    cudaMemcpy(d_st, &st, sizeof(Settings), cudaMemcpyHostToDevice);
    Geom* d_geoms = upload(geoms);
    Material* d_mats = upload(mats);

    if (mode == "raygen") {
        const int iter = atoi(argv[3]);
        const int P = cam.resolution.x * cam.resolution.y;
        PathSegment* d_paths = nullptr;
        cudaMalloc(&d_paths, sizeof(PathSegment) * P);
        const dim3 bs(8, 8), grid((cam.resolution.x + 7) / 8, (cam.resolution.y + 7) / 8);   // pathtrace.cu:430-433
        generateRayFromCamera<<<grid, bs>>>(cam, iter, depth, d_paths, d_st);
        cudaDeviceSynchronize();
        write_file(argv[4], d_paths, (size_t)P);
    } else if (mode == "intersect") {
        std::vector<PathSegment> segs = read_file<PathSegment>(argv[3]);
        const int n = (int)segs.size();
        PathSegment* d_paths = upload(segs);
        ShadeableIntersection* d_is = nullptr;
        cudaMalloc(&d_is, sizeof(ShadeableIntersection) * n);
        cudaMemset(d_is, 0, sizeof(ShadeableIntersection) * n);                              // pathtrace.cu:466
        computeIntersections<<<(n + 63) / 64, 64>>>(0, n, d_paths, d_geoms, (int)geoms.size(), d_is, d_tris, nullptr,
                                                    d_bvh, d_st);
        cudaDeviceSynchronize();
        write_file(argv[4], d_is, (size_t)n);
    } else if (mode == "shade") {
        const int iter = atoi(argv[3]);
        std::vector<PathSegment> segs = read_file<PathSegment>(argv[4]);
        std::vector<ShadeableIntersection> is = read_file<ShadeableIntersection>(argv[5]);
        const int n = (int)segs.size();
        PathSegment* d_paths = upload(segs);
        ShadeableIntersection* d_is = upload(is);
        shadeMaterials<<<(n + 63) / 64, 64>>>(iter, n, d_st, d_is, d_paths, d_mats, d_textures);
        cudaDeviceSynchronize();
        write_file(argv[6], d_paths, (size_t)n);
    } else {
        fprintf(stderr, "unknown mode\n");
        return 1;
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { fprintf(stderr, "CUDA error: %s\n", cudaGetErrorString(e)); return 2; }
    return 0;
}
