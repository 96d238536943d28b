/* b200_pt.h -- C ABI of the B200-native wavefront path tracer.
 *
 * Drop-in boundary for the reference's render driver, path_tracer/src/pathtrace.h:6-9:
 *
 *   void InitDataContainer(GuiDataContainer*)        -> b200pt_set_options
 *   void pathtraceInit(Scene*)                       -> b200pt_create (+ b200pt_reset)
 *   void pathtraceFree()                             -> b200pt_destroy
 *   void pathtrace(uchar4* pbo, int frame, int iter) -> b200pt_render (+ b200pt_tonemap_u8,
 *                                                       b200pt_read_sum)
 *
 * The POD structs below have the SAME byte layout as the reference's sceneStructs.h types
 * (Geom 272 B, Material 48 B, Camera 84 B: path_tracer/src/sceneStructs.h:24-72, offsets in
 * SURVEY.md section 8), so a maintainer's shim can pass scene->geoms.data() etc. unchanged.
 * Matrices are column-major float[16] exactly as glm::mat4 stores them.
 *
 * Conventions kept from the reference: `iteration` is 1-based and is both an RNG seed
 * component and the divisor of the 8-bit preview (pathtrace.cu:57-62,65-86, main.cpp:148-151);
 * the accumulation image is the float SUM over iterations, index x + y*W, 3 floats per pixel
 * (pathtrace.cu:347-356,524); per-(iteration, index, depth) seeding reproduces the reference's
 * random paths bit for bit.
 * Differences: errors are returned (0 = ok, else cudaError_t or B200_E_*), nothing exits;
 * several renderers may coexist (state lives in the handle, not in file statics); iterations
 * are rendered in batches without any per-bounce host round trip.  No CPU fallback.
 *
 * Plain C: pointers and sizes only (a stream is a void*).
 */
#ifndef B200_PT_H
#define B200_PT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { B200PT_SPHERE = 0, B200PT_CUBE = 1, B200PT_MESH = 2 }; /* GeomType, sceneStructs.h:12-17 */

/* == struct Geom (272 bytes) */
typedef struct b200pt_geom {
    int32_t type;
    int32_t materialid;
    float translation[3];
    float rotation[3];
    float scale[3];
    float transform[16];
    float inverseTransform[16];
    float invTranspose[16];
    int32_t tri_start_idx;
    int32_t tri_end_idx;
    int32_t boundingbox_idx;
    float min_bound[3];
    float max_bound[3];
} b200pt_geom;

/* == struct Material (48 bytes) */
typedef struct b200pt_material {
    float color[3];
    float specular_exponent;
    float specular_color[3];
    float hasReflective;
    float hasRefractive;
    float indexOfRefraction;
    float emittance;
    int32_t texture_id; /* -1 = none, else an index into b200pt_scene_ext::textures */
} b200pt_material;

/* == struct Camera (84 bytes) */
typedef struct b200pt_camera {
    int32_t resolution[2];
    float position[3];
    float lookAt[3];
    float view[3];
    float up[3];
    float right[3];
    float fov[2];
    float pixelLength[2];
} b200pt_camera;

typedef struct b200pt_scene {
    b200pt_camera camera;
    int32_t trace_depth;          /* RenderState::traceDepth */
    int32_t n_geoms;
    const b200pt_geom* geoms;
    int32_t n_materials;
    const b200pt_material* materials;
} b200pt_scene;

/* ---- triangle meshes, BVH and textures (the room.json path of the reference) -----------------
 * Same byte layouts as the reference's host structures, so a shim passes scene->triangles.data()
 * and scene->bvh_tree as they are:
 *   b200pt_triangle == class Triangle (124 B, sceneStructs.h:103-160): WORLD-space vertices, per-vertex
 *     texture coordinates and normals, `id` = index of the triangle before the BVH build reordered
 *     the array (Geom::tri_start_idx/tri_end_idx are ranges of ids);
 *   b200pt_bvh_node == class BVHTree (40 B, BVH_tree.h:57-64): flattened in DFS order, the left child
 *     of node i is node i+1, `sub_areas` > 0 marks a leaf holding triangles
 *     [first_area_idx, first_area_idx + sub_areas).
 * A texture is width x height x components bytes, row-major, top row first (class Texture,
 * sceneStructs.h:162-189; only components == 3 yields a colour, anything else reads as black). */
typedef struct b200pt_triangle {
    int32_t id;
    float vertices[3][3];
    float uv[3][2];
    float normals[3][3];
    float bbox_min[3];            /* unused (never assigned by the reference's loader) */
    float bbox_max[3];
} b200pt_triangle;

typedef struct b200pt_bvh_node {
    float min_corner[3];
    float max_corner[3];
    int32_t sub_areas;
    int32_t axis;
    int32_t first_area_idx;
    int32_t rchild_idx;
} b200pt_bvh_node;

typedef struct b200pt_texture {
    int32_t width, height, components;
    const unsigned char* pixels;  /* host memory */
} b200pt_texture;

typedef struct b200pt_scene_ext {
    int32_t n_triangles;
    const b200pt_triangle* triangles;
    int32_t n_bvh_nodes;
    const b200pt_bvh_node* bvh;
    int32_t n_textures;
    const b200pt_texture* textures;
} b200pt_scene_ext;

/* The run-time toggles of GuiDataContainer (utilities.h:17-34) with the reference's defaults,
 * plus the knobs that are new here. */
typedef struct b200pt_options {
    int32_t russian_roulette;   /* default 1 */
    int32_t sort_by_material;   /* default 0 */
    int32_t ssaa;               /* default 1 */
    int32_t dof;                /* default 1 */
    float aperture;             /* default 0.1 */
    float focal_len;            /* default 10.0 */
    int32_t first_bounce_cache; /* NEW, default 0; only honoured when ssaa == 0 && dof == 0 (the camera
                                   rays must not change between iterations); silently off otherwise */
    int32_t batch;              /* NEW: iterations traced concurrently in one wavefront
                                   (0 = library default).  1 keeps the reference's accumulation
                                   order so the float sum image is bit-identical to it.       */
    int32_t geom_bvh;           /* NEW: closest hit over the sphere / cube primitives through a bounding
                                   volume hierarchy instead of the reference's linear scan
                                   (pathtrace.cu:247-290).  It only prunes primitives a ray cannot touch:
                                   the hit found -- hence the image -- is bit-identical either way.
                                   0 = automatic (on), 1 = always, -1 = never (the reference's loop). */
    int32_t mesh_brute_force;   /* !GuiDataContainer::useBVHtree (utilities.h:24; default 0 = walk the BVH).
                                   1: every MESH primitive runs the reference's meshIntersectionTest
                                   (intersections.cu:120-167, pathtrace.cu:273-275) -- all triangles at positions
                                   [tri_start_idx, tri_end_idx), its own (non-barycentric, sic) normal / uv blend.
                                   The reference's GUI toggle for it is compiled out (preview.cpp:247-253); a host
                                   that sets the field gets the reference's result for it, bit for bit. */
    int32_t mesh_no_bbox;       /* !GuiDataContainer::useBBox (utilities.h:25; default 0): with mesh_brute_force,
                                   skip the slab test against Geom::min_bound / max_bound */
    int32_t use_thrust_partition; /* GuiDataContainer::useThrustPartition (utilities.h:27; default 0).  The reference
                                   removes terminated paths either with its own mark / scan / scatter
                                   (pathtrace.cu:377-407) or with thrust::stable_partition (:416-420,499-503); both are
                                   a STABLE live-first partition with the same result.  This library has one
                                   implementation of exactly that (single-pass look-back); the flag is accepted
                                   and recorded so that a host which sets it gets the reference's image for it. */
} b200pt_options;

typedef struct b200pt b200pt;

typedef struct b200pt_stats {
    uint64_t path_segments;   /* sum over bounces of live paths entering intersection */
    uint64_t iterations;      /* iterations accumulated since the last reset                     */
    uint64_t kernel_launches; /* kernels launched since the last reset                           */
    float last_render_ms;     /* CUDA-event time of the most recent b200pt_render call           */
    int32_t traced_depth;     /* GuiDataContainer::TracedDepth of the last iteration             */
    double bounce_kernel_ms;  /* with profiling on: summed device time of the k_bounce launches  */
    uint64_t bounce_kernel_launches; /* ... and how many launches that covers                    */
} b200pt_stats;

void b200pt_default_options(b200pt_options* o);

/* Allocate device state for `scene` on the current device and zero the accumulation image. */
int b200pt_create(const b200pt_scene* scene, b200pt** out);
/* The same for a scene with MESH geoms and / or textured materials (`ext` may be NULL). */
int b200pt_create_ext(const b200pt_scene* scene, const b200pt_scene_ext* ext, b200pt** out);
int b200pt_destroy(b200pt* h);
int b200pt_set_options(b200pt* h, const b200pt_options* o);
/* Zero the accumulation image and the statistics (what pathtraceInit does, pathtrace.cu:121-122). */
int b200pt_reset(b200pt* h);

/* Re-upload geometry, materials and camera of a scene with the same counts / resolution / depth
 * (the reference re-runs pathtraceInit on every camera change, main.cpp:142-146).  Host pointers;
 * synchronises `stream`.  The accumulation image is kept. */
int b200pt_update_scene(b200pt* h, const b200pt_scene* scene, void* stream);
/* Per-kernel CUDA-event timing of the bounce kernel (read back through b200pt_get_stats). */
int b200pt_set_profiling(b200pt* h, int enable);

/* Trace iterations first_iter .. first_iter + (n_iters-1)*iter_stride (1-based, stride >= 1) and
 * add them to the accumulation image.  Asynchronous on `stream` unless `sync` != 0. */
int b200pt_render(b200pt* h, int first_iter, int n_iters, int iter_stride, void* stream, int sync);

/* sendImageToPBO (pathtrace.cu:65-86): dev_rgba8[i] = clamp(int(sum/iter*255.0),0,255), w = 0.
 * dev_rgba8 is a DEVICE buffer of W*H*4 bytes. */
int b200pt_tonemap_u8(b200pt* h, int iter_divisor, void* dev_rgba8, void* stream);

/* The same for ANY device float SUM image of this renderer's resolution (3*W*H floats) -- e.g. the all-reduced image
 * of a multi-GPU render (include/b200_mg.h). */
int b200pt_tonemap_buffer_u8(b200pt* h, const void* dev_sum_rgb, int iter_divisor, void* dev_rgba8, void* stream);

/* Copy the float SUM image (3*W*H floats) to host memory; synchronises. */
int b200pt_read_sum(b200pt* h, float* host_rgb);
/* Replace the float SUM image by 3*W*H floats from host memory (restartable rendering: resume from a
 * stored accumulation; the caller continues with the next iteration number).  Synchronises. */
int b200pt_write_sum(b200pt* h, const float* host_rgb);
/* Device pointer of the float SUM image (for NCCL all-reduce across GPUs). */
int b200pt_sum_device_ptr(b200pt* h, void** dev_ptr, size_t* n_floats);

int b200pt_get_stats(b200pt* h, b200pt_stats* s);

/* Per-stage entry points used by the parity tests (tier T1): each runs ONE stage on arrays in the
 * reference's own AoS layouts so the result can be compared bit for bit with the reference's
 * kernel on identical inputs.  All pointers are DEVICE pointers.
 *   segment = struct PathSegment (48 B: origin 3f, direction 3f, colour 3f, pixelIndex,
 *             remainingBounces, bounces), isect = struct ShadeableIntersection (28 B: t, normal 3f,
 *             materialId, texture_coord 2f)                       sceneStructs.h:78-97 */
int b200pt_debug_raygen(b200pt* h, int iter, void* d_segments /* W*H */, void* stream);
int b200pt_debug_intersect(b200pt* h, int n, const void* d_segments, void* d_isects, void* stream);
int b200pt_debug_shade(b200pt* h, int iter, int n, const void* d_isects, void* d_segments /* in/out */,
                       void* stream);

const char* b200pt_error_string(int code);

#ifdef __cplusplus
}
#endif
#endif /* B200_PT_H */
