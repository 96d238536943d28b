// mesh.hpp -- triangle meshes, BVH construction and textures for the host-side scene loader.
//
// Mirrors what the reference does at load time for `"TYPE": "mesh"` objects and `TEXTURE_FILE`
// materials (path_tracer/src/scene.cpp:94-175, BVH_tree.cpp:27-180, sceneStructs.h:162-174), so that
// the triangle array (in its post-build order), the flattened BVH and the texels handed to the
// device are byte-identical to the reference's own (tests/test_host_cli.py compares with dumps
// written by the reference's unmodified loader):
//   * OBJ reading follows tinyobjloader 1.0.6's rules (fan triangulation; its own decimal parser,
//     which is NOT strtod -- see parse_real);
//   * vertices are transformed to world space at load time, normals by the inverse transpose
//     (not re-normalised), texture coordinates copied;
//   * the BVH is the reference's 7-bucket surface-area-heuristic build with its quirks kept
//     (the zero-box test of BoundingBox::operator||, bucket i left out of both sides of split i,
//     std::partition / std::nth_element orderings), flattened depth first.
#pragma once
#include <string>
#include <vector>

#include "../../include/b200_pt.h"
#include "linalg.hpp"

namespace b200host {

struct TextureImage {
    int width = 0, height = 0, components = 0;
    std::vector<unsigned char> pixels;          // row-major, top row first
};

// Appends the triangles of an OBJ file (world space, ids continuing from triangles.size()) and
// returns the mesh's bounds as the reference computes them (max starts at FLT_MIN, sic).
void load_obj_mesh(const std::string& path, const Mat4& transform, const Mat4& inv_transpose,
                   std::vector<b200pt_triangle>& triangles, float min_bound[3], float max_bound[3]);

// Reorders `triangles` (leaf order) and returns the flattened BVH; empty input -> empty BVH.
std::vector<b200pt_bvh_node> build_bvh_tree(std::vector<b200pt_triangle>& triangles);

// Baseline JPEG (jpeg.cpp: the same bytes stb_image returns, which is what the reference loads textures
// with) and binary PPM (P6, maxval 255) / PGM (P5) are decoded natively.  Throws on other formats.
TextureImage load_texture(const std::string& path);
bool looks_like_jpeg(const unsigned char* data, size_t size);
TextureImage decode_jpeg(const unsigned char* data, size_t size, const std::string& name);
// PNG (png.cpp: 1/2/4/8 bits per channel, all colour types, Adam7; components as stb_image reports them)
bool looks_like_png(const unsigned char* data, size_t size);
TextureImage decode_png(const unsigned char* data, size_t size, const std::string& name);

}  // namespace b200host
