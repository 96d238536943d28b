// integration/pathtrace_b200.cpp -- the reference-side binding for the render driver.
//
// This is the file a maintainer of the reference adds INSTEAD OF pathtrace.cu, intersections.cu,
// interactions.cu and stream_compaction/*.cu (path_tracer/src/): it defines the four functions of
// pathtrace.h:6-9 against the reference's own Scene / GuiDataContainer types and forwards to the
// C ABI of include/b200_pt.h.  Geom, Material, Camera, Triangle and BVHTree are passed as they lie in
// memory (the ABI's POD structs have the same byte layout, sceneStructs.h:24-72,120-160,
// BVH_tree.h); textures as (width, height, components, host pixels).  Conventions kept
// (pathtrace.cu:108-177,423-528): file-static state, Free tolerates being called before Init, the
// toggles are read from the GuiDataContainer on every call, the float SUM image is copied back to
// scene->state.image on every call, the 8-bit preview goes to the caller's DEVICE buffer.
//
// oracle/Makefile compiles it against /root/reference together with the reference's unmodified
// scene.cpp / BVH_tree.cpp / utilities.cpp and the headless main into oracle/_ref/dropin_pathtracer;
// tests/test_dropin_gpu.py checks its output is bit-identical to the reference build's.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include <cuda_runtime.h>

#include "pathtrace.h"
#include "scene.h"
#include "utilities.h"

#include "b200_pt.h"

static_assert(sizeof(Geom) == sizeof(b200pt_geom) && sizeof(Material) == sizeof(b200pt_material) &&
                  sizeof(Camera) == sizeof(b200pt_camera) && sizeof(Triangle) == sizeof(b200pt_triangle) &&
                  sizeof(BVHTree) == sizeof(b200pt_bvh_node), "POD layouts must agree");

static Scene* hst_scene = NULL;
static GuiDataContainer* guiData = NULL;
static b200pt* renderer = NULL;
static void* pinned_image = NULL;

static void check(int code, const char* what)
{
    if (code == 0) return;
    fprintf(stderr, "b200pt (%s): %s\n", what, b200pt_error_string(code));
    exit(EXIT_FAILURE);                  // the reference's own error convention (checkCUDAErrorFn)
}

void InitDataContainer(GuiDataContainer* imGuiData) { guiData = imGuiData; }

void pathtraceInit(Scene* scene)
{
    hst_scene = scene;
    b200pt_scene s;
    s.camera = *reinterpret_cast<const b200pt_camera*>(&scene->state.camera);
    s.trace_depth = scene->state.traceDepth;
    s.n_geoms = (int)scene->geoms.size();
    s.geoms = reinterpret_cast<const b200pt_geom*>(scene->geoms.data());
    s.n_materials = (int)scene->materials.size();
    s.materials = reinterpret_cast<const b200pt_material*>(scene->materials.data());
    // triangle meshes, the flattened BVH and the textures of the scene (scene.cpp:59-69,94-173,218)
    std::vector<b200pt_texture> tex;
    for (const Texture& t : scene->textures) tex.push_back({t.width, t.height, t.components, t.image});
    b200pt_scene_ext e;
    e.n_triangles = (int)scene->triangles.size();
    e.triangles = reinterpret_cast<const b200pt_triangle*>(scene->triangles.data());
    e.n_bvh_nodes = scene->triangles.empty() ? 0 : scene->n_bvh_nodes;
    e.bvh = reinterpret_cast<const b200pt_bvh_node*>(scene->bvh_tree);
    e.n_textures = (int)tex.size();
    e.textures = tex.data();
    check(b200pt_create_ext(&s, &e, &renderer), "pathtraceInit");
    // The interface copies the SUM image back on EVERY call (pathtrace.cu:524): page-lock the caller's vector once so
    // that copy runs at PCIe speed (a pageable 7.68 MB copy per iteration would dominate the call).  Best effort.
    pinned_image = scene->state.image.data();
    if (cudaHostRegister(pinned_image, scene->state.image.size() * sizeof(glm::vec3), cudaHostRegisterDefault) != cudaSuccess) {
        cudaGetLastError();
        pinned_image = NULL;
    }
}

void pathtraceFree()
{
    if (pinned_image) { cudaHostUnregister(pinned_image); pinned_image = NULL; }
    b200pt_destroy(renderer);            // no-op on NULL, like cudaFree(NULL) in the reference
    renderer = NULL;
}

void pathtrace(uchar4* pbo, int /*frame*/, int iter)
{
    b200pt_options o;
    b200pt_default_options(&o);
    if (guiData) {
        o.russian_roulette = guiData->russianRoulette;
        o.sort_by_material = guiData->sortbyMaterial;
        o.ssaa = guiData->SSAA;
        o.dof = guiData->DoF;
        o.aperture = guiData->aperture;
        o.focal_len = guiData->focal_len;
        o.mesh_brute_force = !guiData->useBVHtree;   // pathtrace.cu:49-50
        o.mesh_no_bbox = !guiData->useBBox;
        o.use_thrust_partition = guiData->useThrustPartition;   // pathtrace.cu:499-503: same stable partition, same image
    }
    o.batch = 1;                          // one iteration per call, the reference's accumulation order
    check(b200pt_set_options(renderer, &o), "options");
    check(b200pt_render(renderer, iter, 1, 1, NULL, 0), "pathtrace");
    if (pbo) check(b200pt_tonemap_u8(renderer, iter, pbo, NULL), "sendImageToPBO");
    check(b200pt_read_sum(renderer, reinterpret_cast<float*>(hst_scene->state.image.data())), "image copy");
    if (guiData) {
        b200pt_stats st;
        check(b200pt_get_stats(renderer, &st), "stats");
        guiData->TracedDepth = st.traced_depth;
    }
}
