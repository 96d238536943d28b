// scene.hpp -- host-side scene description and loader of the B200 path tracer.
//
// Mirrors the reference's `Scene` (path_tracer/src/scene.h:17-36, scene.cpp:16-219): the same JSON
// format in, the same plain-old-data out -- `geoms`, `materials`, `state.camera`,
// `state.iterations`, `state.traceDepth`, `state.imageName`, `state.image` -- with the PODs in the
// byte layout of include/b200_pt.h (== sceneStructs.h), so they can be handed to b200pt_create
// as they lie in memory.  Extensions, all optional and inert for files written
// for the reference: material keys REFRACTIVE / IOR (the reference's loader never sets
// hasRefractive / indexOfRefraction, F11).
//
// `"TYPE": "mesh"` objects (OBJ files under <scene dir>/Models) and `TEXTURE_FILE` materials
// (<scene dir>/Textures) are loaded by mesh.cpp; the BVH over all triangles is built at the end of
// loading, as scene.cpp:218 does.
#pragma once
#include <string>
#include <vector>

#include "../../include/b200_pt.h"
#include "mesh.hpp"

namespace b200host {

struct RenderState {                 // sceneStructs.h:69-76
    b200pt_camera camera;
    unsigned int iterations;
    int traceDepth;
    std::vector<float> image;        // 3 floats per pixel, index x + y*W: the float SUM image
    std::string imageName;
};

class Scene {
public:
    // Throws std::runtime_error on unreadable / malformed files and unsupported content.
    explicit Scene(const std::string& filename);

    std::vector<b200pt_geom> geoms;
    std::vector<b200pt_material> materials;
    std::vector<std::string> materialNames;   // id -> name (ids follow the alphabetical order of the names)
    std::vector<b200pt_triangle> triangles;   // world space, in the order the BVH build leaves them
    std::vector<b200pt_bvh_node> bvh_tree;    // flattened depth first (BVH_tree.cpp:168-180)
    std::vector<TextureImage> textures;
    RenderState state;

    // The camera the reference actually renders with: main.cpp re-derives position / view / up /
    // right from spherical coordinates before the first iteration (main.cpp:59-73,121-134; `UP`
    // from the file is ignored, `right` is left un-normalised, the sign of view.x is lost in phi).
    void applyInteractiveCameraSetup();

    b200pt_scene pod() const;        // the view b200pt_create / b200pt_update_scene take
    // triangles / BVH / textures for b200pt_create_ext; the view points into this Scene
    b200pt_scene_ext pod_ext() const;

    // "B2PTSCN1" dump (header, Camera, Geom[], Material[]), "B2PTSCN2" when there are triangles or
    // textures (+ counts, Triangle[], BVHTree[], textures): the fixture format of tests/golden
    void dump(const std::string& path) const;

private:
    mutable std::vector<b200pt_texture> texture_views_;
    void loadFromJSON(const std::string& jsonName);
};

}  // namespace b200host
