// scene.cpp -- the JSON scene loader (see scene.hpp).  Every float computed here follows the
// operation order of the reference's loader so that the PODs are byte-identical to its own
// (tests/test_host_cli.py compares against dumps written by the reference's unmodified scene.cpp).
#include "scene.hpp"

#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>

#include "json_min.hpp"
#include "linalg.hpp"

namespace b200host {

namespace {

const float kPi = 3.1415926535897932384626422832795028841971f;   // utilities.h:12

Vec3 vec3_of(const Json& a)
{
    return {a.at((size_t)0).as_float(), a.at((size_t)1).as_float(), a.at((size_t)2).as_float()};
}
void store(float* dst, const Vec3& v) { dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; }
Vec3 load(const float* p) { return {p[0], p[1], p[2]}; }

// utilityCore::buildTransformationMatrix (utilities.cpp:84-92): T * (Rx * Ry * Rz) * S,
// angles in degrees: rotation.x * (float)PI / 180
Mat4 build_transform(const Vec3& t, const Vec3& r, const Vec3& s)
{
    const Mat4 I = identity();
    const Mat4 translation = translate(I, t);
    Mat4 rot = rotate(I, r.x * kPi / 180, Vec3{1, 0, 0});
    rot = mul(rot, rotate(I, r.y * kPi / 180, Vec3{0, 1, 0}));
    rot = mul(rot, rotate(I, r.z * kPi / 180, Vec3{0, 0, 1}));
    const Mat4 sc = scale(I, s);
    return mul(mul(translation, rot), sc);
}

}  // namespace

Scene::Scene(const std::string& filename)
{
    const size_t dot_at = filename.find_last_of('.');
    if (dot_at == std::string::npos || filename.substr(dot_at) != ".json")
        throw std::runtime_error("Couldn't read from " + filename);      // scene.cpp:26-30
    loadFromJSON(filename);
}

void Scene::loadFromJSON(const std::string& jsonName)
{
    std::ifstream f(jsonName, std::ios::binary);
    if (!f) throw std::runtime_error("cannot open " + jsonName);
    std::stringstream buf;
    buf << f.rdbuf();
    const Json data = Json::parse(buf.str());
    // scene.cpp:37: Models/ and Textures/ live next to the scene file
    const size_t slash = jsonName.find_last_of("/\\");
    const std::string dir = slash == std::string::npos ? std::string(".") : jsonName.substr(0, slash);

    // ---- Materials (scene.cpp:39-72): ids in the alphabetical order of the names -----------------
    std::map<std::string, int> id_of;
    for (const auto& item : data.at("Materials").obj) {
        const Json& p = item.second;
        b200pt_material m;
        std::memset(&m, 0, sizeof(m));                                    // Material newMaterial{}
        Vec3 col{0.f, 0.f, 0.f};
        if (p.contains("RGB")) col = vec3_of(p.at("RGB"));
        const Vec3 spec = p.contains("SPECRGB") ? vec3_of(p.at("SPECRGB")) : col;
        m.specular_exponent = p.value("SPECEX", 1.0f);
        m.hasReflective = p.value("REFLECTIVE", 0.0f);
        m.emittance = p.value("EMITTANCE", 0.0f);
        store(m.color, col);
        store(m.specular_color, spec);
        m.texture_id = -1;
        const std::string tex = p.value("TEXTURE_FILE", std::string());
        if (!tex.empty()) {                                               // scene.cpp:59-69
            m.texture_id = (int)textures.size();
            textures.push_back(load_texture(dir + "/Textures/" + tex));
        }
        // extension keys (ignored by the reference's loader)
        m.hasRefractive = p.value("REFRACTIVE", 0.0f);
        m.indexOfRefraction = p.value("IOR", 0.0f);
        id_of[item.first] = (int)materials.size();
        materialNames.push_back(item.first);
        materials.push_back(m);
    }

    // ---- Objects (scene.cpp:75-175) ------------------------------------------------------------------
    int n_meshes = 0;
    for (const Json& p : data.at("Objects").arr) {
        b200pt_geom g;
        std::memset(&g, 0, sizeof(g));
        const std::string& mat = p.at("MATERIAL").as_string();
        auto it = id_of.find(mat);
        // the reference's std::map::operator[] would silently create id 0 for an unknown name
        g.materialid = it == id_of.end() ? 0 : it->second;
        const Vec3 t = vec3_of(p.at("TRANS")), r = vec3_of(p.at("ROTAT")), s = vec3_of(p.at("SCALE"));
        store(g.translation, t);
        store(g.rotation, r);
        store(g.scale, s);
        const Mat4 xf = build_transform(t, r, s);
        const Mat4 inv = inverse(xf);
        const Mat4 it_ = inverse_transpose(xf);
        std::memcpy(g.transform, xf.m, sizeof(xf.m));
        std::memcpy(g.inverseTransform, inv.m, sizeof(inv.m));
        std::memcpy(g.invTranspose, it_.m, sizeof(it_.m));
        const std::string& type = p.at("TYPE").as_string();
        if (type == "cube") g.type = B200PT_CUBE;
        else if (type == "sphere") g.type = B200PT_SPHERE;
        else if (type == "mesh") {                                        // scene.cpp:94-173
            g.type = B200PT_MESH;
            g.tri_start_idx = (int)triangles.size();
            g.boundingbox_idx = n_meshes++;
            load_obj_mesh(dir + "/Models/" + p.at("OBJ_FILE").as_string(), xf, it_, triangles, g.min_bound, g.max_bound);
            g.tri_end_idx = (int)triangles.size();
        }
        else throw std::runtime_error("unknown object TYPE \"" + type + "\"");
        geoms.push_back(g);
    }

    // ---- Camera (scene.cpp:178-214) -------------------------------------------------------------------
    const Json& cd = data.at("Camera");
    b200pt_camera& cam = state.camera;
    std::memset(&cam, 0, sizeof(cam));
    cam.resolution[0] = cd.at("RES").at((size_t)0).as_int();
    cam.resolution[1] = cd.at("RES").at((size_t)1).as_int();
    const float fovy = cd.at("FOVY").as_float();
    state.iterations = (unsigned)cd.at("ITERATIONS").as_int();
    state.traceDepth = cd.at("DEPTH").as_int();
    state.imageName = cd.at("FILE").as_string();
    const Vec3 pos = vec3_of(cd.at("EYE")), look = vec3_of(cd.at("LOOKAT")), up = vec3_of(cd.at("UP"));
    store(cam.position, pos);
    store(cam.lookAt, look);
    store(cam.up, up);
    const float yscaled = std::tan(fovy * (kPi / 180));
    const float xscaled = (yscaled * cam.resolution[0]) / cam.resolution[1];
    const float fovx = (std::atan(xscaled) * 180) / kPi;
    cam.fov[0] = fovx;
    cam.fov[1] = fovy;
    // scene.cpp:204 computes `right` from the still-zero `view`; applyInteractiveCameraSetup()
    // overwrites it, exactly as main.cpp does before the first iteration
    store(cam.right, normalize(cross(Vec3{0.f, 0.f, 0.f}, up)));
    cam.pixelLength[0] = 2 * xscaled / (float)cam.resolution[0];
    cam.pixelLength[1] = 2 * yscaled / (float)cam.resolution[1];
    store(cam.view, normalize(look - pos));
    state.image.assign(3 * (size_t)cam.resolution[0] * cam.resolution[1], 0.0f);
    bvh_tree = build_bvh_tree(triangles);                                 // scene.cpp:218
}

void Scene::applyInteractiveCameraSetup()
{
    b200pt_camera& cam = state.camera;
    // main.cpp:59-73
    const Vec3 view = load(cam.view);
    const Vec3 viewXZ{view.x, 0.0f, view.z}, viewZY{0.0f, view.y, view.z};
    const float phi = std::acos(dot(normalize(viewXZ), Vec3{0, 0, -1}));
    const float theta = std::acos(dot(normalize(viewZY), Vec3{0, 1, 0}));
    const Vec3 look = load(cam.lookAt);
    const float zoom = length(load(cam.position) - look);
    // main.cpp:121-134
    Vec3 p;
    p.x = zoom * std::sin(phi) * std::sin(theta);
    p.y = zoom * std::cos(theta);
    p.z = zoom * std::cos(phi) * std::sin(theta);
    const Vec3 v = -normalize(p);
    const Vec3 r = cross(v, Vec3{0, 1, 0});
    store(cam.view, v);
    store(cam.up, cross(r, v));
    store(cam.right, r);
    store(cam.position, p + look);
}

b200pt_scene Scene::pod() const
{
    b200pt_scene s;
    s.camera = state.camera;
    s.trace_depth = state.traceDepth;
    s.n_geoms = (int)geoms.size();
    s.geoms = geoms.data();
    s.n_materials = (int)materials.size();
    s.materials = materials.data();
    return s;
}

b200pt_scene_ext Scene::pod_ext() const
{
    texture_views_.clear();
    for (const TextureImage& t : textures) texture_views_.push_back({t.width, t.height, t.components, t.pixels.data()});
    b200pt_scene_ext e;
    e.n_triangles = (int)triangles.size();
    e.triangles = triangles.data();
    e.n_bvh_nodes = (int)bvh_tree.size();
    e.bvh = bvh_tree.data();
    e.n_textures = (int)texture_views_.size();
    e.textures = texture_views_.data();
    return e;
}

void Scene::dump(const std::string& path) const
{
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    const int hdr[4] = {(int)geoms.size(), (int)materials.size(), state.traceDepth, (int)state.iterations};
    const bool v2 = !triangles.empty() || !textures.empty();
    std::fwrite(v2 ? "B2PTSCN2" : "B2PTSCN1", 1, 8, f);
    std::fwrite(hdr, sizeof(int), 4, f);
    std::fwrite(&state.camera, sizeof(b200pt_camera), 1, f);
    std::fwrite(geoms.data(), sizeof(b200pt_geom), geoms.size(), f);
    std::fwrite(materials.data(), sizeof(b200pt_material), materials.size(), f);
    if (v2) {
        const int ext[4] = {(int)triangles.size(), (int)bvh_tree.size(), (int)textures.size(), 0};
        std::fwrite(ext, sizeof(int), 4, f);
        std::fwrite(triangles.data(), sizeof(b200pt_triangle), triangles.size(), f);
        std::fwrite(bvh_tree.data(), sizeof(b200pt_bvh_node), bvh_tree.size(), f);
        for (const TextureImage& t : textures) {
            const int th[4] = {t.width, t.height, t.components, 0};
            std::fwrite(th, sizeof(int), 4, f);
            std::fwrite(t.pixels.data(), 1, t.pixels.size(), f);
        }
    }
    std::fclose(f);
}

}  // namespace b200host
