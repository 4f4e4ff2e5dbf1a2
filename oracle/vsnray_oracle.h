/*
 * vsnray_oracle.h -- TEST INFRASTRUCTURE ONLY.  Plain-C CPU restatement of the reference's
 * ray-query hot path (tu500/visionaray @ 175062fa).  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library, and only as the
 * checker or the reported CPU baseline -- never as part of the product path.
 *
 * Parity status: PINNED.  The restatement is checked bit-for-bit (hit, ids, t, u, v,
 * primitive_list_index) against the reference's own headers compiled here
 * (oracle/_ref/libvsnray_ref.so, -ffp-contract=off) by tests/test_oracle_vs_ref.py, against
 * the reference's known-answer test test/unittests/get_normal.cpp:17-75, and against golden
 * vectors generated from the reference (tests/golden/, generator tests/golden/make_golden.py).
 *
 * Each function cites the reference file:line it follows (paths relative to
 * /root/reference/include/visionaray/).
 */
#ifndef VSNRAY_ORACLE_H
#define VSNRAY_ORACLE_H 1

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Same layouts as include/vsnray_b200.h (kept separate so the oracle has no dependency on
 * the product headers). */
typedef struct { float ori[3]; float pad0; float dir[3]; float pad1; } vso_ray;
typedef struct { uint32_t geom_id, prim_id, pad0[2]; float v1[3], pad1, e1[3], pad2, e2[3], pad3; } vso_triangle;
typedef struct { uint32_t geom_id, prim_id, pad0[2]; float center[3], pad1, radius, pad2[3]; } vso_sphere;
/* generic_primitive<basic_triangle<3,float>, basic_sphere<float>> (generic_primitive.h, variant.h): the
 * alternative stored in place (64 bytes), then its 1-based index (1 triangle, 2 sphere); 80 bytes, 16-aligned */
typedef struct { uint32_t storage[16]; uint32_t which; uint32_t pad[3]; } vso_generic_primitive;
typedef struct { float bbox_min[3]; uint32_t first; float bbox_max[3]; uint32_t num_prims; } vso_bvh_node;
typedef struct { int32_t hit; uint32_t prim_id, geom_id; float t, u, v; uint32_t prim_list_index, pad; } vso_hit;
/* kind 0 plastic, 1 matte (as before); 2 emissive: cd = ce, kd = ls (material.h emissive<T>);
 * 3 mirror: cs = cr, ks = kr, ca = ior, cd = absorption (material.h mirror<T>, brdf.h:160-212) */
typedef struct { float ca[3], ka, cd[3], kd, cs[3], ks, spec_exp; int32_t kind; float pad[2]; } vso_material;
/* kind 0 point_light, 1 spot_light (detail/spot_light.inl:14-38) */
typedef struct { float pos[3], kl, cl[3], att_const, att_lin, att_quad; int32_t kind; float spot_exponent, spot_dir[3], spot_cos_cutoff; } vso_point_light;

typedef struct {
    int32_t  kernel;          /* 0 simple, 1 whitted, 2 ao, 3 pathtracing */
    int32_t  pixel_sampler;   /* 0 uniform, 1 jittered, 2 jittered_blend, 3/4/5 ssaa 2x/4x/8x */
    int32_t  width, height;
    float    eye[4], center[4], up[4];
    float    fovy, aspect;
    uint32_t frame_num, seed;
    uint32_t num_bounces;
    float    epsilon;
    float    bg[4];
    float    ambient[4];
    int32_t  ao_samples;
    float    ao_radius;
    int32_t  n_lights, n_materials;
} vso_frame_params;

typedef struct {
    int prim_kind;                 /* 0 triangles, 1 spheres, 2 generic_primitive<triangle, sphere> */
    const void* prims; size_t n_prims;
    const vso_bvh_node* nodes; size_t n_nodes;
    const uint32_t* indices; size_t n_indices;   /* NULL: primitive(i) = prims[i] (bvh.h:177-180) */
} vso_bvh;

/* per-ray-batch visit counters for the roofline denominator:
 * counters[0] += inner-node steps (each reads both 32-B children), counters[1] += leaf
 * primitives tested, counters[2] += rays */
typedef struct { uint64_t n_inner, n_prim, n_rays, n_hits; } vso_counters;

/* primitives (math/intersect.h) */
void vso_intersect_triangle(const vso_ray* r, const vso_triangle* tri, vso_hit* out);
void vso_intersect_sphere(const vso_ray* r, const vso_sphere* s, vso_hit* out);
void vso_intersect_generic(const vso_ray* r, const vso_generic_primitive* g, vso_hit* out);
int  vso_intersect_aabb(const vso_ray* r, const float bmin[3], const float bmax[3],
                        const float inv_dir[3], float* tnear, float* tfar);

/* intersect<ClosestHit|AnyHit>(ray, bvh, isect, max_t) -- detail/bvh/intersect.inl:35-134 */
void vso_intersect_bvh(int any, const vso_ray* r, const vso_bvh* b, float max_t, vso_hit* out,
                       vso_counters* c);

/* closest_hit()/any_hit() over a ray list (detail/traverse_linear.inl:186-318, one bvh) */
void vso_query(int any, const vso_bvh* b, const vso_ray* rays, size_t n_rays, float max_t,
               const float* max_t_per_ray, vso_hit* hits, vso_counters* counters, int nthreads);

/* multi_hit<N>() over a ray list and a range of bvhs (detail/traverse_linear.inl:325-380,
 * detail/multi_hit.h): hits[i * n + k] = k-th closest record of ray i, misses (t = FLT_MAX) behind */
void vso_multi_hit_bvh(const vso_ray* r, const vso_bvh* b, int n, vso_hit* result);
void vso_multi_hit(const vso_bvh* bvhs, int n_bvhs, const vso_ray* rays, size_t n_rays, int n, vso_hit* hits, int nthreads);

/* brute force over all primitives in list order (detail/traverse_linear.inl:26-63); used to
 * classify t-ties */
void vso_query_brute(int any, int prim_kind, const void* prims, size_t n_prims,
                     const vso_ray* rays, size_t n_rays, float max_t, vso_hit* hits, int nthreads);

/* camera basis (detail/cuda_sched.inl:259-266) and primary ray (detail/sched_common.h:130-150) */
void vso_camera_basis(const vso_frame_params* p, float eye[3], float u[3], float v[3], float w[3]);
void vso_primary_ray(float x, float y, int width, int height, const float eye[3],
                     const float cu[3], const float cv[3], const float cw[3], vso_ray* out);

/* deterministic counter-based sampler shared by oracle, reference harness and CUDA kernels */
uint32_t vso_sampler_init(uint32_t pixel, uint32_t frame, uint32_t seed);
float    vso_sampler_next(uint32_t* state);

/* frame(): per pixel ray-gen + kernel + store/blend (detail/sched_common.h:361-528 and
 * detail/{simple,whitted,pathtracing}.inl, src/examples/ao/main.cpp:183-246).
 * rgba: width*height*4 floats, row-major, y up.  counters may be NULL. */
void vso_frame(const vso_frame_params* p, const vso_bvh* b, const float* normals16,
               const vso_material* mats, const vso_point_light* lights, float* rgba,
               vso_counters* counters, int nthreads);

/* matrix camera (sched_params with view_matrix / proj_matrix, scheduler.h:60-96) and the depth
 * buffer it enables.  Matrices are the reference's matrix<4,4,float>: 16 floats, column-major.
 *   vso_mat4_inverse      -- math/detail/matrix4.inl:209-244
 *   vso_primary_ray_matrix -- detail/sched_common.h:152-178
 *   vso_depth_transform    -- detail/sched_common.h:337-354
 *   vso_frame_matrix       -- vso_frame() with that camera; depth (width*height floats) may be
 *                             NULL.  depth = hit ? transform(isect_pos) : 1 (:425), stored /
 *                             blended / SSAA-accumulated like the colour (:559-575, :713-765) */
void  vso_mat4_inverse(const float m[16], float out[16]);
void  vso_primary_ray_matrix(float x, float y, int width, int height, const float inv_view[16],
                             const float inv_proj[16], vso_ray* out);
float vso_depth_transform(const float pos[3], const float view[16], const float proj[16]);
void  vso_frame_matrix(const vso_frame_params* p, const float view[16], const float proj[16],
                       const vso_bvh* b, const float* normals16, const vso_material* mats,
                       const vso_point_light* lights, float* rgba, float* depth,
                       vso_counters* counters, int nthreads);

/* get_surface() inputs beyond per-face normals (kernels.h:194-319, get_surface.h:330-530).
 * texture_ref<vector<4, unorm<8>>, 2> (texture/detail/texture2d.h): RGBA8 texels, row-major. */
typedef struct {
    const void* data; int32_t width, height;
    int32_t filter;         /* 0 Nearest, 1 Linear (texture/forward.h tex_filter_mode) */
    int32_t address[2];     /* 0 Wrap, 1 Mirror, 2 Clamp (tex_address_mode; texture/detail/filter/common.h:24-48) */
} vso_texture;
typedef struct {
    int32_t normal_binding;      /* 2 = the params carry NO normals: get_normal(hr, primitive) -- triangle
                                    normalize(cross(e1, e2)), sphere (isect_pos - center) / radius (get_normal.h:96-118);
                                    0 normals_per_face_binding: normals[prim_id]; 1 normals_per_vertex_binding:
                                    shading normal = normalize(lerp(normals[3 prim_id + 0..2], u, v))
                                    (get_shading_normal.h:60-76), geometric normal = normalize(cross(e1, e2)) of
                                    primitive(primitive_list_index) (get_surface.h:285-328, get_normal.h:96-101) */
    int32_t color_binding;       /* 0 none, 1 colors_per_face_binding, 2 colors_per_vertex_binding (get_color.h:24-58) */
    const float* normals;        /* vector<3,float>: 16 B each */
    const float* colors;         /* vector<3,float>: 16 B each */
    const float* tex_coords;     /* vector<2,float>: 8 B each, 3 per primitive (get_tex_coord.h:24-40); NULL = none */
    const vso_texture* textures; /* by geom_id; width or height 0 -> tex_color 1 (get_surface.h:443-445) */
    int32_t n_textures;
} vso_surface;

/* tex2D(texture_ref<vector<4, unorm<8>>, 2>, coord) -- texture/detail/sampler2d.h:81-113,
 * filter/{nearest,linear}.h; out = 4 floats */
void vso_tex2d(const vso_texture* tex, float s, float t, float out[4]);

/* vso_frame() with kernel_params that carry a surface description and a
 * generic_material<emissive, matte, mirror, plastic> list (surf may be NULL = vso_frame) */
void vso_frame_surface(const vso_frame_params* p, const vso_bvh* b, const vso_surface* surf,
                       const vso_material* mats, const vso_point_light* lights, float* rgba,
                       vso_counters* counters, int nthreads);

/* Colour formats of the render target (pixel_format.h): what pixel_access::store / get / blend convert to
 * and from (detail/pixel_access.h:38-58,668-689,1153-1175; detail/color_conversion.h:168-265;
 * math/detail/unorm.inl:18-31): 0 PF_RGBA32F (16 B), 1 PF_RGBA8 (4 B), 2 PF_RGB32F (16 B: vector<3,float> is 16-byte aligned, last word
 * unused), 3 PF_RGB8 (3 B).
 * vso_frame_format(): vso_frame() into a buffer of that format (blend and SSAA accumulate THROUGH the format). */
void vso_store_pixel(int format, const float rgba[4], void* dst);
void vso_get_pixel(int format, const void* src, float rgba[4]);
void vso_frame_format(const vso_frame_params* p, const vso_bvh* b, const float* normals16,
                      const vso_material* mats, const vso_point_light* lights, void* color, int format,
                      vso_counters* counters, int nthreads);

/* max depth of the BVH2 (root = depth 1); the reference's traversal stack holds 31 entries
 * (detail/stack.h:39-47, detail/bvh/intersect.inl:60) */
int vso_bvh_max_depth(const vso_bvh* b);

#ifdef __cplusplus
}
#endif
#endif
