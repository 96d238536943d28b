// TEST / BASELINE INFRASTRUCTURE -- not product code.
//
// Headless driver for the UNMODIFIED reference CUDA path tracer.  It is compiled
// (by oracle/Makefile) together with the reference's own pathtrace.cu,
// intersections.cu, interactions.cu, scene.cpp, BVH_tree.cpp, utilities.cpp and
// stream_compaction/*.cu straight from /root/reference; only the reference's
// GL/GLFW/ImGui translation units (main.cpp, preview.cpp, glslUtility.cpp,
// ImGui/*) are replaced by this file, because the box has no graphics stack.
//
// What it restates from the reference's main.cpp (and nothing else):
//   * the spherical-coordinate camera re-derivation that runCuda() performs
//     before the first iteration            (path_tracer/src/main.cpp:59-73,121-134)
//   * the iteration loop: ++iteration; pathtrace(pbo, 0, iteration)
//                                            (path_tracer/src/main.cpp:142-160)
// The PBO becomes a plain cudaMalloc'd uchar4 buffer.
//
// Output: the float SUM image (3*W*H floats, index x + y*W, exactly
// scene->state.image as the reference leaves it) written raw to --out, and one
// JSON line with wall time on stdout.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "scene.h"
#include "pathtrace.h"
#include "utilities.h"
#include "image.h"

// The reference declares but never defines the destructor (scene.h:24).
Scene::~Scene() {}

// saveImage() of the reference (path_tracer/src/main.cpp:88-112), its loop restated on the reference's OWN Image
// class (image.cpp:22-44 is compiled in unmodified): mean = sum / samples, x mirrored, clamp * 255 truncated, PNG
// through stb_image_write.  `base` replaces the time-stamped name.  Used to pin host/image_io.cpp to the reference.
static void ref_save_image(const glm::vec3* sum, int width, int height, float samples, const std::string& base, bool hdr)
{
    Image img(width, height);
    for (int x = 0; x < width; x++)
        for (int y = 0; y < height; y++) {
            int index = x + (y * width);
            glm::vec3 pix = sum[index];
            img.setPixel(width - 1 - x, y, glm::vec3(pix) / samples);
        }
    img.savePNG(base);
    if (hdr) img.saveHDR(base);
}

static void usage(const char* a0)
{
    fprintf(stderr,
        "usage: %s SCENE.json [--iters N] [--out sum.bin] [--pbo pbo.bin] [--save-png BASE] [--sort] [--thrust-partition]\n"
        "       %s --convert SUM.bin W H SAMPLES BASE [--hdr]   (the reference's saveImage on a stored SUM image)\n"
        "       [--no-rr] [--no-ssaa] [--no-dof] [--aperture A] [--focal F]\n"
        "       [--refract MATERIAL_ID IOR]...  (patches Material after load: the reference\n"
        "        loader has no key for hasRefractive/indexOfRefraction)\n", a0, a0);
}

int main(int argc, char** argv)
{
    if (argc < 2) { usage(argv[0]); return 1; }
    if (std::string(argv[1]) == "--convert") {      // SUM.bin W H SAMPLES BASE [--hdr]: saveImage only, no GPU involved
        if (argc < 7) { usage(argv[0]); return 1; }
        const int w = atoi(argv[3]), h = atoi(argv[4]);
        std::vector<glm::vec3> sum((size_t)w * h);
        FILE* f = fopen(argv[2], "rb");
        if (!f || fread(sum.data(), sizeof(glm::vec3), sum.size(), f) != sum.size()) { fprintf(stderr, "cannot read %s\n", argv[2]); return 1; }
        fclose(f);
        ref_save_image(sum.data(), w, h, (float)atof(argv[5]), argv[6], argc > 7 && std::string(argv[7]) == "--hdr");
        return 0;
    }
    const char* sceneFile = argv[1];
    int iters = -1;
    std::string out, pbo_out, save_png;
    GuiDataContainer* gui = new GuiDataContainer();
    std::vector<std::pair<int, float>> refract;
    for (int i = 2; i < argc; ++i) {
        std::string a = argv[i];
        if (a == "--iters" && i + 1 < argc) iters = atoi(argv[++i]);
        else if (a == "--out" && i + 1 < argc) out = argv[++i];
        else if (a == "--pbo" && i + 1 < argc) pbo_out = argv[++i];
        else if (a == "--save-png" && i + 1 < argc) save_png = argv[++i];
        else if (a == "--sort") gui->sortbyMaterial = true;
        else if (a == "--thrust-partition") gui->useThrustPartition = true;
        else if (a == "--no-rr") gui->russianRoulette = false;
        else if (a == "--no-bvh") gui->useBVHtree = false;
        else if (a == "--no-bbox") gui->useBBox = false;
        else if (a == "--no-ssaa") gui->SSAA = false;
        else if (a == "--no-dof") gui->DoF = false;
        else if (a == "--aperture" && i + 1 < argc) gui->aperture = (float)atof(argv[++i]);
        else if (a == "--focal" && i + 1 < argc) gui->focal_len = (float)atof(argv[++i]);
        else if (a == "--refract" && i + 2 < argc) {
            int id = atoi(argv[++i]); float ior = (float)atof(argv[++i]);
            refract.push_back({id, ior});
        } else { usage(argv[0]); return 1; }
    }

    Scene* scene = new Scene(sceneFile);   // leaked on purpose, as the reference does
    for (auto& r : refract) {
        if (r.first < 0 || r.first >= (int)scene->materials.size()) { fprintf(stderr, "bad material id\n"); return 1; }
        scene->materials[r.first].hasRefractive = 1.0f;
        scene->materials[r.first].indexOfRefraction = r.second;
    }
    RenderState* rs = &scene->state;
    Camera& cam = rs->camera;
    if (iters < 0) iters = (int)rs->iterations;
    const int width = cam.resolution.x, height = cam.resolution.y;

    // ---- main.cpp:59-73 ----
    glm::vec3 view = cam.view;
    glm::vec3 viewXZ = glm::vec3(view.x, 0.0f, view.z);
    glm::vec3 viewZY = glm::vec3(0.0f, view.y, view.z);
    float phi = glm::acos(glm::dot(glm::normalize(viewXZ), glm::vec3(0, 0, -1)));
    float theta = glm::acos(glm::dot(glm::normalize(viewZY), glm::vec3(0, 1, 0)));
    float zoom = glm::length(cam.position - cam.lookAt);
    // ---- main.cpp:121-134 (camchanged is true on the first frame) ----
    glm::vec3 cameraPosition;
    cameraPosition.x = zoom * sin(phi) * sin(theta);
    cameraPosition.y = zoom * cos(theta);
    cameraPosition.z = zoom * cos(phi) * sin(theta);
    cam.view = -glm::normalize(cameraPosition);
    glm::vec3 v = cam.view;
    glm::vec3 u = glm::vec3(0, 1, 0);
    glm::vec3 r = glm::cross(v, u);
    cam.up = glm::cross(r, v);
    cam.right = r;
    cam.position = cameraPosition;
    cameraPosition += cam.lookAt;
    cam.position = cameraPosition;

    if (getenv("REF_DUMP_CAMERA")) {
        printf("{\"camera\": {\"position\": [%.9g, %.9g, %.9g], \"view\": [%.9g, %.9g, %.9g], "
               "\"up\": [%.9g, %.9g, %.9g], \"right\": [%.9g, %.9g, %.9g], \"pixelLength\": [%.9g, %.9g]}}\n",
               cam.position.x, cam.position.y, cam.position.z, cam.view.x, cam.view.y, cam.view.z,
               cam.up.x, cam.up.y, cam.up.z, cam.right.x, cam.right.y, cam.right.z,
               cam.pixelLength.x, cam.pixelLength.y);
    }

    uchar4* pbo = nullptr;
    if (cudaMalloc(&pbo, sizeof(uchar4) * width * height) != cudaSuccess) {
        fprintf(stderr, "cudaMalloc failed (no GPU?)\n");
        return 2;
    }
    InitDataContainer(gui);
    pathtraceFree();          // main.cpp:144 calls Free before the first Init
    pathtraceInit(scene);
    cudaDeviceSynchronize();

    auto t0 = std::chrono::high_resolution_clock::now();
    for (int iteration = 1; iteration <= iters; ++iteration)
        pathtrace(pbo, 0, iteration);
    cudaDeviceSynchronize();
    auto t1 = std::chrono::high_resolution_clock::now();
    double ms = std::chrono::duration<double, std::milli>(t1 - t0).count();

    if (!out.empty()) {
        FILE* f = fopen(out.c_str(), "wb");
        if (!f) { perror("fopen"); return 1; }
        fwrite(rs->image.data(), sizeof(glm::vec3), (size_t)width * height, f);
        fclose(f);
    }
    if (!save_png.empty()) ref_save_image(rs->image.data(), width, height, (float)iters, save_png, false);
    if (!pbo_out.empty()) {
        std::vector<uchar4> h((size_t)width * height);
        cudaMemcpy(h.data(), pbo, h.size() * sizeof(uchar4), cudaMemcpyDeviceToHost);
        FILE* f = fopen(pbo_out.c_str(), "wb");
        if (!f) { perror("fopen"); return 1; }
        fwrite(h.data(), sizeof(uchar4), h.size(), f);
        fclose(f);
    }
    printf("{\"impl\": \"reference_cuda\", \"scene\": \"%s\", \"width\": %d, \"height\": %d, \"depth\": %d, "
           "\"iters\": %d, \"ms_total\": %.3f, \"ms_per_iter\": %.4f, \"spp_per_s\": %.3f}\n",
           sceneFile, width, height, rs->traceDepth, iters, ms, ms / iters, iters / (ms * 1e-3));
    pathtraceFree();
    cudaFree(pbo);
    return 0;
}
