// TEST / BASELINE INFRASTRUCTURE -- not product code.
//
// Dumps what the UNMODIFIED reference scene loader (path_tracer/src/scene.cpp:33-219, compiled
// from /root/reference by oracle/Makefile) produces for a scene file, as the plain-old-data the
// render driver consumes: Camera (84 B), Geom[] (272 B each), Material[] (48 B each).  The camera
// is the one the reference actually renders with, i.e. after the spherical-coordinate
// re-derivation runCuda() performs before the first iteration (main.cpp:59-73,121-134).
// Runs on the CPU (no GPU needed).  Used to generate tests/golden/*.scene.bin
// (tests/golden/make_pt_golden.py) which pin the repo's own scene loader and feed the C ABI.
//
// File layout (little endian): "B2PTSCN1", int32 n_geoms, n_materials, trace_depth, iterations,
// then Camera, Geom[n_geoms], Material[n_materials].  Scenes with triangle meshes are written as
// "B2PTSCN2": the same, followed by int32 n_triangles, n_bvh_nodes, n_textures, 0, then
// Triangle[n_triangles] (124 B each, in the order the BVH build left them), BVHTree[n_bvh_nodes]
// (40 B each, DFS order), and per texture int32 width, height, components, 0 + the raw pixel bytes.
//
//   ref_scene_dump SCENE.json OUT.bin [--refract MATERIAL_ID IOR]...
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "scene.h"
#include "utilities.h"

Scene::~Scene() {}   // declared but never defined by the reference (scene.h:24)

int main(int argc, char** argv)
{
    if (argc < 3) { fprintf(stderr, "usage: %s SCENE.json OUT.bin [--refract ID IOR]...\n", argv[0]); return 1; }
    Scene* scene = new Scene(argv[1]);
    for (int i = 3; i + 2 < argc + 0 && std::string(argv[i]) == "--refract"; i += 3) {
        int id = atoi(argv[i + 1]);
        if (id < 0 || id >= (int)scene->materials.size()) { fprintf(stderr, "bad material id\n"); return 1; }
        scene->materials[id].hasRefractive = 1.0f;
        scene->materials[id].indexOfRefraction = (float)atof(argv[i + 2]);
    }
    Camera& cam = scene->state.camera;

    // ---- main.cpp:59-73 ----
    glm::vec3 view = cam.view;
    glm::vec3 viewXZ = glm::vec3(view.x, 0.0f, view.z);
    glm::vec3 viewZY = glm::vec3(0.0f, view.y, view.z);
    float phi = glm::acos(glm::dot(glm::normalize(viewXZ), glm::vec3(0, 0, -1)));
    float theta = glm::acos(glm::dot(glm::normalize(viewZY), glm::vec3(0, 1, 0)));
    float zoom = glm::length(cam.position - cam.lookAt);
    // ---- main.cpp:121-134 ----
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

    static_assert(sizeof(Camera) == 84 && sizeof(Geom) == 272 && sizeof(Material) == 48, "layout");
    static_assert(sizeof(Triangle) == 124 && sizeof(BVHTree) == 40, "layout");
    const bool v2 = !scene->triangles.empty() || !scene->textures.empty();
    FILE* f = fopen(argv[2], "wb");
    if (!f) { perror("fopen"); return 1; }
    int hdr[4] = {(int)scene->geoms.size(), (int)scene->materials.size(), scene->state.traceDepth,
                  (int)scene->state.iterations};
    fwrite(v2 ? "B2PTSCN2" : "B2PTSCN1", 1, 8, f);
    fwrite(hdr, sizeof(int), 4, f);
    fwrite(&cam, sizeof(Camera), 1, f);
    for (Geom g : scene->geoms) {
        if (g.type != MESH) {            // uninitialised in the reference for non-mesh geoms
            g.tri_start_idx = g.tri_end_idx = g.boundingbox_idx = 0;
            g.min_bound = g.max_bound = glm::vec3(0.0f);
        }
        fwrite(&g, sizeof(Geom), 1, f);
    }
    fwrite(scene->materials.data(), sizeof(Material), scene->materials.size(), f);
    if (v2) {
        int ext[4] = {(int)scene->triangles.size(), scene->triangles.empty() ? 0 : scene->n_bvh_nodes,
                      (int)scene->textures.size(), 0};
        fwrite(ext, sizeof(int), 4, f);
        for (Triangle t : scene->triangles) {
            memset(&t.BoundBox, 0, sizeof(t.BoundBox));     // never assigned by the reference's loader
            fwrite(&t, sizeof(Triangle), 1, f);
        }
        for (int i = 0; i < ext[1]; ++i) {
            BVHTree node = scene->bvh_tree[i];
            if (node.sub_areas > 0) { node.Axis = 0; node.RChild_idx = 0; }        // uninitialised in leaves
            else node.first_area_idx = 0;                                          // ... and in inner nodes
            fwrite(&node, sizeof(BVHTree), 1, f);
        }
        for (const Texture& t : scene->textures) {
            int th[4] = {t.width, t.height, t.components, 0};
            fwrite(th, sizeof(int), 4, f);
            fwrite(t.image, 1, (size_t)t.width * t.height * t.components, f);
        }
    }
    fclose(f);
    printf("wrote %s: %d geoms, %d materials, %dx%d, depth %d, %d iterations\n", argv[2], hdr[0], hdr[1],
           cam.resolution.x, cam.resolution.y, hdr[2], hdr[3]);
    return 0;
}
