// main.cpp -- headless command-line front end of the B200 path tracer.
//
// Drop-in for the reference executable (path_tracer/src/main.cpp:34-86,114-168):
//
//     cuda_path_tracer SCENEFILE.json
//
// loads the scene, renders `Camera.ITERATIONS` samples per pixel and writes
// "<FILE>.<YYYY-MM-DD_HH-MM-SSz>.<N>samp.png" into the current directory (main.cpp:88-112), the
// float SUM image having been accumulated exactly as the reference's pathtrace() loop would.
// The GLFW window / ImGui panel of the reference do not exist here (no graphics stack on the
// box); its nine GUI toggles (utilities.h:17-34) are command-line flags with the same defaults:
//
//   --iters N            override Camera.ITERATIONS
//   --no-rr              russianRoulette = false          --sort        sortbyMaterial = true
//   --no-ssaa            SSAA = false                     --no-dof      DoF = false
//   --thrust-partition   useThrustPartition = true (pathtrace.cu:499-503): the same stable partition, the same image
//   --aperture A         (default 0.1)                    --focal F     focal_len (default 10.0)
//   --first-bounce-cache reuse bounce 0's hits (honoured only with --no-ssaa --no-dof)
//   --no-bvh             useBVHtree = false: meshes through meshIntersectionTest (every triangle of a mesh)
//   --no-bbox            useBBox = false: ... without its bounding-box cull
//   --geom-bvh on|off|auto   closest hit over spheres / cubes through a BVH instead of the linear scan
//                        (same image either way; auto = on)
//   --batch B            iterations in flight (0 = library default, 1 = the reference's exact
//                        accumulation order: bit-identical float SUM image)
//   --iter-first F --iter-stride S   trace iterations F, F+S, ... (N of them): the multi-GPU split
//   --out BASE           write BASE.png instead of the reference's time-stamped name
//   --hdr                also write BASE.hdr              --no-png      skip the PNG
//   --sum-out FILE       raw float SUM image (3*W*H floats, index x + y*W)
//   --pbo-out FILE       raw uchar4 preview buffer (sendImageToPBO, pathtrace.cu:65-86)
//   --dump-scene FILE    write the loaded scene as a "B2PTSCN1"/"B2PTSCN2" POD dump and exit (no GPU needed)
//   --device D           CUDA device ordinal
//   --gpus K             render on CUDA devices 0..K-1 of this box in ONE process (include/b200_mg.h): one host thread per
//                        device, device g traces iterations g+1, g+1+K, ..., one ncclAllReduce of the float SUM image;
//                        the image equals the one-GPU image up to fp32 re-association of the per-pixel sums
//   --preview-every N    every N iterations: (all-reduce,) read the SUM image back and write
//                        "<FILE>.<time>.<iterations so far>samp.png" -- the mid-run save of the reference's 'S' key
//                        (main.cpp:176-182) for a headless run; rendering continues
//   --json               one JSON line with the timing on stdout
//   --resume-sum FILE --samples-done K   restartable rendering: start from a float SUM image that
//                        already holds K samples (a --sum-out of an earlier run); iterations
//                        continue at K+1 and the output is divided by K + N
//   --convert SUM.bin W H SAMPLES BASE   (first argument) only run saveImage on a stored SUM image
// Exit codes follow the reference: 1 on a usage error, -1 (255) when the scene cannot be read.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include <cuda_runtime_api.h>

#include "../../include/b200_mg.h"
#include "../../include/b200_pt.h"
#include "image_io.hpp"
#include "scene.hpp"

using namespace b200host;

namespace {

void usage(const char* argv0) { std::printf("Usage: %s SCENEFILE.json\n", argv0); }

void write_raw(const std::string& path, const void* data, size_t bytes)
{
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f || std::fwrite(data, 1, bytes, f) != bytes) {
        std::fprintf(stderr, "cannot write %s\n", path.c_str());
        std::exit(EXIT_FAILURE);
    }
    std::fclose(f);
}

void check(int code, const char* what)
{
    if (code == 0) return;
    std::fprintf(stderr, "b200pt (%s): %s\n", what, b200pt_error_string(code));
    std::exit(EXIT_FAILURE);              // the reference's own error convention (checkCUDAErrorFn)
}

void mg_check(int code, const char* what)
{
    if (code == 0) return;
    std::fprintf(stderr, "b200mg (%s): %s\n", what, b200mg_error_string(code));
    std::exit(EXIT_FAILURE);
}

}  // namespace

int main(int argc, char** argv)
{
    const std::string startTimeString = currentTimeString();
    if (argc < 2) { usage(argv[0]); return 1; }
    if (std::string(argv[1]) == "--convert") {           // SUM image -> PNG (and HDR), no GPU involved
        if (argc < 7) { usage(argv[0]); return 1; }
        const int w = std::atoi(argv[3]), h = std::atoi(argv[4]);
        const float samples = (float)std::atof(argv[5]);
        std::vector<float> sum(3 * (size_t)w * h);
        FILE* f = std::fopen(argv[2], "rb");
        if (!f || std::fread(sum.data(), sizeof(float), sum.size(), f) != sum.size()) {
            std::fprintf(stderr, "cannot read %zu floats from %s\n", sum.size(), argv[2]);
            return EXIT_FAILURE;
        }
        std::fclose(f);
        const Image img = image_from_sum(sum.data(), w, h, samples);
        img.savePNG(argv[6]);
        if (argc > 7 && std::string(argv[7]) == "--hdr") img.saveHDR(argv[6]);
        return EXIT_SUCCESS;
    }
    const char* sceneFile = argv[1];

    b200pt_options opt;
    b200pt_default_options(&opt);
    int iters = -1, device = 0, iter_first = 1, iter_stride = 1, gpus = 1, preview_every = 0;
    bool json_stats = false;
    bool want_png = true, want_hdr = false;
    int samples_done = 0;
    bool iter_first_given = false;
    std::string out_base, sum_out, pbo_out, dump_scene, resume_sum;
    std::vector<std::pair<int, float>> refract;
    for (int i = 2; i < argc; ++i) {
        const std::string a = argv[i];
        auto next = [&]() -> const char* {
            if (i + 1 >= argc) { usage(argv[0]); std::exit(1); }
            return argv[++i];
        };
        if (a == "--iters") iters = std::atoi(next());
        else if (a == "--no-rr") opt.russian_roulette = 0;
        else if (a == "--sort" || a == "--sort-material") opt.sort_by_material = 1;
        else if (a == "--thrust-partition") opt.use_thrust_partition = 1;   // useThrustPartition: same stable partition, same image
        else if (a == "--no-ssaa") opt.ssaa = 0;
        else if (a == "--no-dof") opt.dof = 0;
        else if (a == "--aperture") opt.aperture = (float)std::atof(next());
        else if (a == "--focal") opt.focal_len = (float)std::atof(next());
        else if (a == "--first-bounce-cache") opt.first_bounce_cache = 1;
        else if (a == "--no-bvh") opt.mesh_brute_force = 1;       // GuiDataContainer::useBVHtree = false
        else if (a == "--no-bbox") opt.mesh_no_bbox = 1;          // GuiDataContainer::useBBox = false
        else if (a == "--batch") opt.batch = std::atoi(next());
        else if (a == "--geom-bvh") {
            const std::string v = next();
            if (v == "on") opt.geom_bvh = 1;
            else if (v == "off") opt.geom_bvh = -1;
            else if (v == "auto") opt.geom_bvh = 0;
            else { usage(argv[0]); return 1; }
        }
        else if (a == "--iter-first") { iter_first = std::atoi(next()); iter_first_given = true; }
        else if (a == "--resume-sum") resume_sum = next();
        else if (a == "--samples-done") samples_done = std::atoi(next());
        else if (a == "--iter-stride") iter_stride = std::atoi(next());
        else if (a == "--out") out_base = next();
        else if (a == "--hdr") want_hdr = true;
        else if (a == "--no-png") want_png = false;
        else if (a == "--sum-out") sum_out = next();
        else if (a == "--pbo-out" || a == "--pbo") pbo_out = next();
        else if (a == "--dump-scene") dump_scene = next();
        else if (a == "--device") device = std::atoi(next());
        else if (a == "--gpus") gpus = std::atoi(next());
        else if (a == "--preview-every") preview_every = std::atoi(next());
        else if (a == "--json") json_stats = true;
        else if (a == "--refract") {      // patch a material after load, as the test tooling does
            const int id = std::atoi(next());
            refract.emplace_back(id, (float)std::atof(next()));
        } else { usage(argv[0]); return 1; }
    }

    // Load scene file (main.cpp:46-47, scene.cpp:16-31)
    std::cout << "Reading scene from " << sceneFile << " ..." << std::endl;
    std::cout << " " << std::endl;
    Scene* scene = nullptr;
    try {
        scene = new Scene(sceneFile);
    } catch (const std::exception& e) {
        std::cout << e.what() << std::endl;
        return -1;
    }
    for (const auto& r : refract) {
        if (r.first < 0 || r.first >= (int)scene->materials.size()) { std::fprintf(stderr, "bad material id\n"); return 1; }
        scene->materials[r.first].hasRefractive = 1.0f;
        scene->materials[r.first].indexOfRefraction = r.second;
    }
    scene->applyInteractiveCameraSetup();             // main.cpp:59-73,121-134 (first frame: camchanged)
    if (!dump_scene.empty()) {
        scene->dump(dump_scene);
        return 0;
    }
    RenderState& state = scene->state;
    const int width = state.camera.resolution[0], height = state.camera.resolution[1];
    const int n_iters = iters >= 0 ? iters : (int)state.iterations;

    // ---- render: one GPU through libb200pt.so, or --gpus K through the multi-GPU layer (libb200mg.so: one host thread
    // per device, iterations split with stride K, ONE ncclAllReduce of the float SUM image per read-back) ----
    auto save_preview = [&](int samples_so_far) {       // what the reference's 'S' key does mid-run (main.cpp:176-182, saveImage)
        const Image img = image_from_sum(state.image.data(), width, height, (float)samples_so_far);
        img.savePNG(out_base.empty() ? output_basename(state.imageName, startTimeString, (float)samples_so_far)
                                     : out_base + "." + std::to_string(samples_so_far) + "samp");
    };
    if (!resume_sum.empty()) {
        FILE* f = std::fopen(resume_sum.c_str(), "rb");
        if (!f || std::fread(state.image.data(), sizeof(float), state.image.size(), f) != state.image.size()) {
            std::fprintf(stderr, "cannot read %zu floats from %s\n", state.image.size(), resume_sum.c_str());
            return EXIT_FAILURE;
        }
        std::fclose(f);
        if (!iter_first_given) iter_first = samples_done + 1;
    }
    const b200pt_scene pod = scene->pod();
    const b200pt_scene_ext ext = scene->pod_ext();
    const int chunk = preview_every > 0 ? preview_every : (n_iters > 0 ? n_iters : 1);
    b200pt_stats st;
    double secs = 0.0;
    float allreduce_ms = 0.f;
    std::vector<unsigned char> pbo;
    if (gpus > 1) {
        if (iter_stride != 1) { std::fprintf(stderr, "--iter-stride and --gpus exclude each other (--gpus splits the iterations itself)\n"); return 1; }
        b200mg* group = nullptr;
        b200mg_pt* renderer = nullptr;
        mg_check(b200mg_create(gpus, nullptr, &group), "device group");
        mg_check(b200mg_pt_create(group, &pod, &ext, &renderer), "pathtraceInit");
        mg_check(b200mg_pt_set_options(renderer, &opt), "options");
        if (!resume_sum.empty()) mg_check(b200mg_pt_write_sum(renderer, state.image.data()), "resume");
        const auto t0 = std::chrono::steady_clock::now();
        for (int done = 0; done < n_iters || done == 0; done += chunk) {
            const int n = std::min(chunk, n_iters - done);
            mg_check(b200mg_pt_render(renderer, iter_first + done, n), "pathtrace");
            mg_check(b200mg_pt_reduce_read_sum(renderer, state.image.data(), &allreduce_ms), "all-reduce + image copy");
            if (preview_every > 0 && done + n < n_iters) save_preview(samples_done + done + n);
            if (n_iters == 0) break;
        }
        secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        mg_check(b200mg_pt_get_stats(renderer, &st), "stats");
        if (!pbo_out.empty()) {
            pbo.resize((size_t)width * height * 4);
            mg_check(b200mg_pt_preview_u8(renderer, iter_first + (n_iters > 0 ? n_iters - 1 : 0), pbo.data()), "sendImageToPBO");
        }
        b200mg_pt_destroy(renderer);                    // pathtraceFree
        b200mg_destroy(group);
    } else {
        if (cudaSetDevice(device) != cudaSuccess) {
            std::fprintf(stderr, "no usable CUDA device %d (this renderer has no CPU fallback)\n", device);
            return EXIT_FAILURE;
        }
        b200pt* renderer = nullptr;
        check(b200pt_create_ext(&pod, &ext, &renderer), "pathtraceInit");
        check(b200pt_set_options(renderer, &opt), "options");
        if (!resume_sum.empty()) check(b200pt_write_sum(renderer, state.image.data()), "resume");
        const auto t0 = std::chrono::steady_clock::now();
        for (int done = 0; done < n_iters || done == 0; done += chunk) {
            const int n = std::min(chunk, n_iters - done);
            check(b200pt_render(renderer, iter_first + done * iter_stride, n, iter_stride, nullptr, 1), "pathtrace");
            check(b200pt_read_sum(renderer, state.image.data()), "image copy");     // pathtrace.cu:524
            if (preview_every > 0 && done + n < n_iters) save_preview(samples_done + done + n);
            if (n_iters == 0) break;
        }
        secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        check(b200pt_get_stats(renderer, &st), "stats");
        if (!pbo_out.empty()) {
            const int last_iter = iter_first + (n_iters > 0 ? n_iters - 1 : 0) * iter_stride;
            void* d_pbo = nullptr;
            if (cudaMalloc(&d_pbo, (size_t)width * height * 4) != cudaSuccess) return EXIT_FAILURE;
            check(b200pt_tonemap_u8(renderer, last_iter, d_pbo, nullptr), "sendImageToPBO");
            pbo.resize((size_t)width * height * 4);
            if (cudaMemcpy(pbo.data(), d_pbo, pbo.size(), cudaMemcpyDeviceToHost) != cudaSuccess) return EXIT_FAILURE;
            cudaFree(d_pbo);
        }
        b200pt_destroy(renderer);                       // pathtraceFree
    }
    std::printf("%d iterations of %dx%d, depth %d: %.3f s, %.1f spp/s, %.1f Mpath-segments/s (%llu segments, %llu launches)\n",
                n_iters, width, height, state.traceDepth, secs, n_iters / secs, (double)st.path_segments / secs / 1e6,
                (unsigned long long)st.path_segments, (unsigned long long)st.kernel_launches);
    if (json_stats)
        std::printf("{\"impl\": \"b200_cli\", \"gpus\": %d, \"iters\": %d, \"width\": %d, \"height\": %d, \"depth\": %d, \"seconds\": %.6f, "
                    "\"spp_per_s\": %.3f, \"mpath_segments_per_s\": %.3f, \"path_segments\": %llu, \"launches\": %llu, "
                    "\"last_allreduce_ms\": %.4f}\n",
                    gpus, n_iters, width, height, state.traceDepth, secs, n_iters / secs, (double)st.path_segments / secs / 1e6,
                    (unsigned long long)st.path_segments, (unsigned long long)st.kernel_launches, allreduce_ms);
    if (!pbo_out.empty()) write_raw(pbo_out, pbo.data(), pbo.size());
    if (!sum_out.empty()) write_raw(sum_out, state.image.data(), state.image.size() * sizeof(float));

    // saveImage (main.cpp:88-112): `samples` is the number of iterations accumulated
    const float samples = (float)(samples_done + n_iters);
    if (samples > 0 && (want_png || want_hdr)) {
        const Image img = image_from_sum(state.image.data(), width, height, samples);
        const std::string base = out_base.empty() ? output_basename(state.imageName, startTimeString, samples) : out_base;
        if (want_png) img.savePNG(base);
        if (want_hdr) img.saveHDR(base);
    }
    return EXIT_SUCCESS;
}
