/*
 * vsnray_b200.h -- C ABI of libvsnray_b200.so: the B200 (sm_100a) ray-query hot path.
 *
 * The reference (tu500/visionaray) has no FFI; its hot path is C++ templates instantiated in
 * the caller's translation unit.  The entry points below are what the host adaptor headers
 * (include/visionaray_b200/) bind in place of those template instantiations.  Each one cites the
 * reference interface it replaces (paths relative to /root/reference/include/visionaray/).
 *
 * Conventions: plain pointers and sizes only; every call returns 0 (VSB_OK) or a negative
 * vsb_* code / positive cudaError_t; vsb_strerror() describes either.  There is no CPU
 * fallback: without a CUDA device every compute entry point fails with VSB_ERR_NO_DEVICE.
 * Calls on one context are serialised by the caller (the reference's frame() is not
 * re-entrant either, detail/cuda_sched.inl:307-320).
 */
#ifndef VSNRAY_B200_H
#define VSNRAY_B200_H 1

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- data layouts (the reference's memory layouts are the ABI) ---- */

/* basic_ray<float>: math/ray.h; vec3 is 16-byte aligned (math/detail/vector3f.inl:12). 32 B. */
typedef struct vsb_ray {
    float ori[3]; float pad0;
    float dir[3]; float pad1;
} vsb_ray;

/* basic_triangle<3,float,unsigned>: math/triangle.h + math/primitive.h. 64 B. */
typedef struct vsb_triangle {
    uint32_t geom_id; uint32_t prim_id; uint32_t pad0[2];
    float v1[3]; float pad1;
    float e1[3]; float pad2;
    float e2[3]; float pad3;
} vsb_triangle;

/* basic_sphere<float,unsigned>: math/sphere.h. 48 B. */
typedef struct vsb_sphere {
    uint32_t geom_id; uint32_t prim_id; uint32_t pad0[2];
    float center[3]; float pad1;
    float radius; float pad2[3];
} vsb_sphere;

/* generic_primitive<basic_triangle<3,float>, basic_sphere<float>> (generic_primitive.h, variant.h): the
 * alternative stored in place, then its 1-based index (1 triangle, 2 sphere).  80 B, 16-byte aligned.
 * A leaf of such a tree mixes both kinds; intersect() dispatches on `which` (detail/generic_primitive.inl:20-44). */
typedef struct vsb_generic_primitive {
    uint32_t storage[16];      /* a vsb_triangle, or a vsb_sphere in the first 48 bytes */
    uint32_t which;
    uint32_t pad[3];
} vsb_generic_primitive;

/* bvh_node: bvh.h:52-119. inner <=> num_prims == 0; children at first, first+1. 32 B. */
typedef struct vsb_bvh_node {
    float    bbox_min[3];
    uint32_t first;        /* first_child (inner) | first_prim (leaf, index into indices[]) */
    float    bbox_max[3];
    uint32_t num_prims;
} vsb_bvh_node;

/* hit_record_bvh<ray, BVH, hit_record<ray, primitive<unsigned>>>:
 * math/intersect.h:88-115 + detail/bvh/hit_record.h:30-46 (isect_pos is not produced by the
 * reference's intersect() either, math/intersect.h:257).  A miss has hit=0, t=FLT_MAX. 32 B. */
typedef struct vsb_hit {
    int32_t  hit;
    uint32_t prim_id;
    uint32_t geom_id;
    float    t;
    float    u;
    float    v;
    uint32_t prim_list_index;   /* primitive_list_index: position in indices[] */
    uint32_t pad;
} vsb_hit;

/* One member of generic_material<emissive, matte, mirror, plastic> (generic_material.h, material.h,
 * detail/material/*.inl), flattened; the adaptor fills it through the reference's getters. 64 B.
 *   VSB_MAT_PLASTIC  ca ka cd kd cs ks spec_exp        VSB_MAT_MATTE    ca ka cd kd
 *   VSB_MAT_EMISSIVE cd = ce, kd = ls                  VSB_MAT_MIRROR   cs = cr, ks = kr, ca = ior, cd = absorption */
typedef struct vsb_material {
    float ca[3]; float ka;
    float cd[3]; float kd;
    float cs[3]; float ks;
    float spec_exp; int32_t kind;   /* VSB_MAT_PLASTIC | MATTE | EMISSIVE | MIRROR (the last two need generic_materials) */
    float pad[2];
} vsb_material;

/* point_light<float> (point_light.h:58-64) or spot_light<float> (spot_light.h, detail/spot_light.inl:14-38),
 * flattened; a light list is homogeneous in the reference (the Lights iterator type) but may mix here. 64 B.
 * Spot: intensity = cl kl att spot, spot = dot(spot_dir, normalize(pos - position)) > spot_cos_cutoff ?
 * pow(spot, spot_exponent) : 0; spot_dir must be a unit vector. */
enum { VSB_LIGHT_POINT = 0, VSB_LIGHT_SPOT = 1 };
typedef struct vsb_point_light {
    float pos[3]; float kl;
    float cl[3];  float att_const;
    float att_lin; float att_quad; int32_t kind; float spot_exponent;
    float spot_dir[3]; float spot_cos_cutoff;
} vsb_point_light;

enum { VSB_MAT_PLASTIC = 0, VSB_MAT_MATTE = 1, VSB_MAT_EMISSIVE = 2, VSB_MAT_MIRROR = 3 };

/* texture_ref<vector<4, unorm<8>>, 2> (texture/detail/texture2d.h, texture_common.h:36-92): RGBA8
 * texels, row-major, texel(x, y) = data[y * width + x].  width or height 0 = "no texture" (tex_color 1,
 * get_surface.h:443-445).  filter: tex_filter_mode Nearest / Linear; address: tex_address_mode per axis
 * (texture/forward.h:18-34). */
enum { VSB_TEX_NEAREST = 0, VSB_TEX_LINEAR = 1, VSB_TEX_BSPLINE = 2, VSB_TEX_CARDINAL_SPLINE = 4 };   /* tex_filter_mode; 3 =
     BSplineInterpol needs the reference's host prefilter (texture/detail/prefilter.h) and is not supported */
enum { VSB_TEX_WRAP = 0, VSB_TEX_MIRROR = 1, VSB_TEX_CLAMP = 2, VSB_TEX_BORDER = 3 };   /* Border falls through to Clamp in the
                                                                                          reference too (filter/common.h:41-46) */
typedef struct vsb_texture {
    const void* data;
    int32_t width, height;
    int32_t filter;
    int32_t address[2];
} vsb_texture;

/* What get_surface() reads besides the hit record: the members of kernel_params beyond prims /
 * materials / lights (kernels.h:194-319) with their binding tags (tags.h).  vector<3,float> is
 * 16 bytes, vector<2,float> 8. */
enum { VSB_NORMALS_PER_FACE = 0, VSB_NORMALS_PER_VERTEX = 1, VSB_NORMALS_FROM_PRIMITIVE = 2 };
enum { VSB_COLORS_NONE = 0, VSB_COLORS_PER_FACE = 1, VSB_COLORS_PER_VERTEX = 2 };
typedef struct vsb_surface_desc {
    int32_t normal_binding;       /* per face: normals[prim_id] (get_normal.h:28-37); per vertex: shading normal =
                                     normalize(lerp(normals[3 prim_id + 0..2], u, v)) (get_shading_normal.h:60-76),
                                     geometric normal = normalize(cross(e1, e2)) (get_normal.h:96-101); from primitive =
                                     kernel_params WITHOUT normals (the default make_kernel_params overload, kernels.h:328-366):
                                     get_normal(hr, primitive) -- triangle normalize(cross(e1, e2)), sphere
                                     (isect_pos - center) / radius (get_normal.h:96-118); `normals` may be NULL.  The
                                     only binding of a sphere scene (frames: simple / whitted kernels) */
    int32_t color_binding;        /* colors[prim_id] or lerp(colors[3 prim_id + 0..2], u, v) (get_color.h:24-58) */
    const float* normals;         /* n_prims (per face) or 3 n_prims (per vertex) x 16 B */
    const float* colors;          /* like normals; NULL with VSB_COLORS_NONE */
    const float* tex_coords;      /* 3 n_prims x 8 B (get_tex_coord.h:24-40); NULL = params without textures */
    const vsb_texture* textures;  /* n_textures entries, indexed by geom_id (geom_id >= n_textures: no texture) */
    int32_t n_textures;
} vsb_surface_desc;
enum { VSB_PRIM_TRIANGLE = 0, VSB_PRIM_SPHERE = 1, VSB_PRIM_GENERIC = 2 };
enum { VSB_KERNEL_SIMPLE = 0, VSB_KERNEL_WHITTED = 1, VSB_KERNEL_AO = 2, VSB_KERNEL_PATHTRACING = 3 };
enum { VSB_PS_UNIFORM = 0, VSB_PS_JITTERED = 1, VSB_PS_JITTERED_BLEND = 2,
       VSB_PS_SSAA2 = 3, VSB_PS_SSAA4 = 4, VSB_PS_SSAA8 = 5 };   /* ssaa_type<2|4|8>, detail/sched_common.h:222-295,640-698 */

/* colour format of the render target = what pixel_access::store / get / blend convert to and from
 * (pixel_format.h, pixel_traits.h, detail/pixel_access.h:38-58,668-689,1153-1175, detail/color_conversion.h:168-265,
 * math/detail/unorm.inl:18-31).  Bytes per pixel: 16, 4, 16 (vector<3,float> is 16-byte aligned in the reference;
 * the 4th word is never written), 3.  unorm8 = (uint32)(clamp(f, 0, 1) * 255.0) (truncation); RGBA -> RGB multiplies
 * by alpha, RGB -> RGBA (blend, SSAA) reads alpha 1; blends and SSAA accumulate THROUGH the format like the reference. */
enum { VSB_PF_RGBA32F = 0, VSB_PF_RGBA8 = 1, VSB_PF_RGB32F = 2, VSB_PF_RGB8 = 3 };

/* traversal engine: the wide quantised BVH8 (default, fast) or the reference-order BVH2 walk
 * (bit-exact with detail/bvh/intersect.inl:35-134 including tie order and any-hit winner). */
enum { VSB_TRAVERSE_WIDE = 0, VSB_TRAVERSE_EXACT = 1,
       /* opt-in, vsb_closest_hit / vsb_any_hit on triangle scenes: the BVH8 walk with the hit / miss decision of the
        * watertight test of Woop, Benthin, Wald (JCGT 2013) instead of the reference's Moeller-Trumbore inside test
        * (math/intersect.h:122-179, which lets rays through on shared edges); t / u / v in the reference's form, so
        * records equal VSB_TRAVERSE_WIDE's wherever both tests accept.  Not the default: the default reproduces the
        * reference's decisions.  (csrc/watertight.cuh; DESIGN.md 4.11 counts the rays that change.) */
       VSB_TRAVERSE_WATERTIGHT = 2 };

/* where the caller's buffers live */
enum { VSB_MEM_DEVICE = 0, VSB_MEM_HOST = 1 };
enum { VSB_TILES_COMPACT = 0, VSB_TILES_IN_PLACE = 1 };

enum {
    VSB_OK = 0,
    VSB_ERR_NO_DEVICE   = -1,
    VSB_ERR_BAD_ARG     = -2,
    VSB_ERR_BAD_TREE    = -3,
    VSB_ERR_NOT_INIT    = -4,
    VSB_ERR_UNSUPPORTED = -5,
    VSB_ERR_TOO_LARGE   = -6        /* a size beyond a stated fixed maximum (vsb_closest_hit_static) */
};

/* Kernel + scheduler parameters of one frame() call.  Mirrors what cuda_sched<R>::frame()
 * passes to render<<<>>> (detail/cuda_sched.inl:60-99,259-266) plus the fields of
 * kernel_params (kernels.h:94-319) the built-in kernels read.  The first block is shared
 * field-for-field with the oracle (oracle/vsnray_oracle.h: vso_frame_params). */
typedef struct vsb_frame_params {
    int32_t  kernel;          /* VSB_KERNEL_* */
    int32_t  pixel_sampler;   /* VSB_PS_* (sched_common.h:29-52) */
    int32_t  width, height;
    float    eye[4], center[4], up[4];   /* pinhole camera (camera.h) */
    float    fovy, aspect;               /* fovy in radians */
    uint32_t frame_num, seed;
    uint32_t num_bounces;
    float    epsilon;
    float    bg[4];
    float    ambient[4];
    int32_t  ao_samples;
    float    ao_radius;
    int32_t  n_lights, n_materials;
} vsb_frame_params;

enum { VSB_CAM_BASIS = 0, VSB_CAM_MATRIX = 1 };

typedef struct vsb_frame_desc {
    vsb_frame_params p;
    /* camera basis as cuda_sched_impl_frame computes it on the host
     * (detail/cuda_sched.inl:259-266); filled by vsb_camera_basis() */
    float    cam_eye[4], cam_u[4], cam_v[4], cam_w[4];
    int32_t  scissor[4];                 /* x, y, w, h; (0,0,width,height) = whole image */
    const float*           normals;      /* per-face normals, 16 B each, indexed by prim_id
                                            (normals_per_face_binding, get_normal.h:28-37); NULL = the
                                            normals attached to scenes[0] by vsb_scene_set_normals() */
    const vsb_material*    materials;    /* indexed by geom_id */
    const vsb_point_light* lights;
    void*    color;                      /* buffer[y*width+x] (pixel_access.h:38-58), pixels of `color_format` (PF_RGBA32F by default) */
    void*    depth;                      /* PF_DEPTH32F, buffer[y*width+x], same memory (and tile layout) as color;
                                            may be NULL.  depth = hit ? ((proj*view*isect_pos).z/w + 1)/2 : 1
                                            (depth_transform, detail/sched_common.h:337-354,425); needs
                                            camera_kind = VSB_CAM_MATRIX like the reference, whose depth
                                            transform only exists for matrix sched_params */
    int32_t  mem;                        /* VSB_MEM_*: where color / ray_count live (the render target) */
    int32_t  traverse;                   /* VSB_TRAVERSE_* */
    /* image-space tile partition for multi-GPU (64x64 tiles, tile k -> rank k % nranks);
     * nranks <= 1 renders everything.  With nranks > 1 `color` is the rank's compact tile
     * buffer (tiles_of_rank * 64*64 pixels), see vsb_tiles_of_rank(). */
    int32_t  tile_rank, tile_nranks;
    uint64_t* ray_count;                 /* optional counter (same memory as color): += rays traced (device),
                                            = rays traced this frame (host) */
    int32_t  params_mem;                 /* VSB_MEM_*: where materials / lights (and normals, if given) live */
    int32_t  tile_target;                /* with tile_nranks > 1: VSB_TILES_COMPACT (default) or VSB_TILES_IN_PLACE =
                                            `color` is the whole row-major frame (width*height pixels), e.g. a peer
                                            GPU's buffer opened with vsb_ipc_open(): the rank's tiles are stored
                                            straight into it over NVLink, no gather pass (device memory, no blend).
                                            With mem = VSB_MEM_HOST the frame must be page-locked and mapped
                                            (cudaHostAlloc / cudaHostRegister with cudaHostRegisterMapped): the
                                            kernel stores finished pixels straight into it, so the ranks of a
                                            node assemble ONE host frame in shared memory */
    /* sched_params built from matrices instead of a camera (make_sched_params(view, proj, rt),
     * scheduler.h:60-96,205-242): rays come from the inverse matrices (detail/sched_common.h:152-178;
     * the inverses are computed per frame on the host like detail/cuda_sched.inl:231-234).
     * matrix<4,4,float> layout: 16 floats, column-major. */
    int32_t  camera_kind;                /* VSB_CAM_BASIS (cam_eye/u/v/w above) or VSB_CAM_MATRIX */
    float    view_matrix[16], proj_matrix[16];
    int32_t  color_format;               /* VSB_PF_*: format of `color` (default VSB_PF_RGBA32F).  Compact multi-rank tile
                                            buffers (VSB_TILES_COMPACT with tile_nranks > 1) are RGBA32F only. */
    int32_t  generic_materials;          /* non-zero: `materials` is a generic_material<emissive, matte, mirror, plastic>
                                            list, i.e. entries may be VSB_MAT_EMISSIVE / VSB_MAT_MIRROR (pathtracing.inl:93-104,
                                            whitted.inl:66-122).  Detected automatically when params_mem == VSB_MEM_HOST. */
    int32_t  spot_lights;                /* non-zero: `lights` may hold VSB_LIGHT_SPOT entries.  Detected automatically when
                                            params_mem == VSB_MEM_HOST. */
    void*    mirror_color;               /* optional, with compact tile buffers (tile_nranks > 1, VSB_TILES_COMPACT), RGBA32F, device
                                            memory: every finished pixel is ALSO stored at its row-major position in this whole
                                            frame (e.g. a peer GPU's buffer opened with vsb_ipc_open()).  Progressive blending
                                            (jittered_blend) keeps its state in the rank's compact buffer while the assembled frame
                                            appears in place over NVLink: no all-gather / de-interleave pass */
} vsb_frame_desc;

typedef struct vsb_scene vsb_scene;

/* ---- context ----
 * One context per process and GPU.  Streams and threads: every query / frame call takes a `cuda_stream`
 * (NULL = the library's own non-blocking stream, see vsb_sync).  The calls of a context share device-side
 * state (scene table, work counters, staging buffers), so the library orders them on the device itself: the
 * stream of a call waits for the work of the previous call, whichever stream that ran on.  Mixing streams is
 * therefore safe but does not overlap two calls.  Calls with device buffers return asynchronously; order the
 * consumers of their outputs on the same stream (or synchronise it).  Host THREADS must serialise their calls
 * into a context, like callers of the reference's frame() (detail/cuda_sched.inl:307-320, not re-entrant). */
int  vsb_init(int device);                       /* cudaSetDevice + stream/queue state */
int  vsb_shutdown(void);
int  vsb_sync(void);                             /* wait for the library's own stream (calls made with cuda_stream == NULL)
                                                    and for the last query / frame call, whichever stream it ran on */
const char* vsb_strerror(int code);
int  vsb_version(void);
int  vsb_device_sm_count(int* out);
/* scheduling knobs of the persistent traversal kernel: "query_service_min" (default 8), "frame_service_min"
 * (0 = per kernel: 32 simple / Whitted, 24 AO / path tracing): idle lanes of a warp before finished rays are
 * serviced / refilled, 1..32; "phase_min" (node/primitive phase vote threshold of the round-1 loop);
 * "trace_queue" (1: primitive tests shared by the warp, the default), "drain_min" (24) / "pend_min" (8) (its
 * batching thresholds); "build_algorithm" and the other builder knobs: see vsb_scene_build.  Results do not
 * depend on the scheduling knobs. */
int  vsb_tune(const char* key, int value);

/* ---- scene: replaces cuda_index_bvh<P>(index_bvh<P> const&) upload (bvh.h:344-350,446-447)
 * and the .ref() it hands to kernels (bvh.h:364-376).  indices == NULL means a non-index
 * bvh (bvh_t, primitive(i) = prims[i], bvh.h:177-180).  The tree is validated, uploaded and
 * re-encoded on the GPU into the wide quantised BVH8. ---- */
int  vsb_scene_create(int prim_kind, const void* prims, size_t n_prims,
                      const vsb_bvh_node* nodes, size_t n_nodes,
                      const uint32_t* indices, size_t n_indices,
                      vsb_scene** out);
/* The same scene with the BVH BUILT ON THE GPU from the primitives alone: replaces
 * build<index_bvh<P>>(first, last) (bvh.h:455 -> detail/bvh/build.inl:28-140 -> binned_sah_builder,
 * detail/bvh/sah.h:150-762; one host thread, recursive) followed by the upload above.
 *   vsb_tune("build_algorithm", 1) (default): THE REFERENCE'S TREE -- the same binned-SAH decisions, partition order
 *     and node numbering, evaluated breadth-first (csrc/build_sah.cuh): vsb_scene_download_bvh returns the nodes and
 *     indices the host builder returns (bounds equal as floats; the sign of a zero may differ), so every result,
 *     any-hit winners and primitive_list_index included, is that of a scene made from the host builder's tree.
 *     No spatial splits (build them with the host builder and pass the tree to vsb_scene_create).
 *   vsb_tune("build_algorithm", 0): Morton sort + parallel locally-ordered clustering + SAH leaf collapse (<= 4
 *     primitives per leaf; csrc/build_ploc.cuh).  Another valid BVH: closest hits equal up to t-ties, any-hit
 *     winners and primitive_list_index follow THAT tree's order.  Builds in about the same time (8.7 vs 13 ms at 5 M
 *     primitives, 54 vs 49 ms at 20 M), traces 5-14 % slower.  Knobs: "build_radius" (1..32, default 8), "build_cost_node_x100", "build_cost_prim_x100".
 * Both lay the tree out in the reference's node / indices format and re-encode it into the BVH8 like any other
 * tree.  prims: host or device memory (mem = VSB_MEM_*). */
int  vsb_scene_build(int prim_kind, const void* prims, size_t n_prims, int mem, vsb_scene** out);
/* the scene's reference-layout tree back on the host -- what index_bvh<P>::nodes() / indices() hold
 * (bvh.h:298-403); sizes from vsb_scene_stats (out[5] nodes; indices = n_prims).  Either pointer may be NULL. */
int  vsb_scene_download_bvh(const vsb_scene* s, vsb_bvh_node* nodes_out, size_t nodes_cap,
                            uint32_t* indices_out, size_t indices_cap);
/* device time of the last vsb_scene_build(): ms4[0] total (events on the library stream, host round
 * trips included), [1] bounds + sort, [2] clustering, [3] layout + wide re-encode; iterations may be NULL */
int  vsb_last_build_ms(float* ms4, int* iterations);
int  vsb_scene_destroy(vsb_scene* s);
/* per-face normals kept on the device with the scene (what a cuda_sched user holds in a thrust
 * device_vector and hands to make_kernel_params, src/viewer/viewer.cpp:507-521); n = n_prims, or more:
 * normals[prim_id] is ONE array over all bvhs of a frame's range, so when a frame runs over several scenes
 * the array attached to scenes[0] (or passed in vsb_frame_desc.normals) must reach the largest prim_id of
 * all of them (vsb_frame returns VSB_ERR_BAD_ARG otherwise) */
int  vsb_scene_set_normals(vsb_scene* s, const float* normals16, size_t n, int mem);
/* The surface description of the scene's triangles, deep-copied to the device (arrays and texels;
 * mem = where desc's arrays live) and used by every later vsb_frame() on this scene in place of the
 * per-face normals: what a cuda_sched user uploads once into thrust vectors / cuda_texture objects and
 * hands to make_kernel_params(normal_binding[, color_binding], ..., normals, tex_coords, materials[, colors],
 * textures, ...) (kernels.h:395-509, src/viewer/viewer.cpp:507-541,797-851).  desc == NULL detaches it.
 * The AO kernel (src/examples/ao) takes per-face normals only. */
int  vsb_scene_set_surface(vsb_scene* s, const vsb_surface_desc* desc, int mem);
/* diagnostics: out[0]=wide nodes, out[1]=wide node bytes, out[2]=leaf prim records,
 * out[3]=prim record bytes, out[4]=bvh2 max depth, out[5]=bvh2 nodes */
int  vsb_scene_stats(const vsb_scene* s, uint64_t* out, int n);
/* copy the re-encoded wide nodes back to the host (property tests of the re-encode) */
int  vsb_scene_download_wide(const vsb_scene* s, void* nodes_out, size_t node_bytes_cap,
                             void* prims_out, size_t prim_bytes_cap);

/* ---- ray queries over a list of rays: replaces closest_hit()/any_hit() over
 * [prims.begin, prims.end) (detail/traverse_linear.inl:186-318) -> intersect<Traversal>()
 * (detail/bvh/intersect.inl:35-134).  scenes[0..n_scenes) mirrors the range of bvh refs.
 * max_t_per_ray may be NULL (then max_t applies to all rays; FLT_MAX = unbounded). ---- */
int  vsb_closest_hit(const vsb_scene* const* scenes, int n_scenes,
                     const vsb_ray* rays, size_t n_rays, float max_t, const float* max_t_per_ray,
                     vsb_hit* hits, int mem, int traverse, void* cuda_stream);
int  vsb_any_hit(const vsb_scene* const* scenes, int n_scenes,
                 const vsb_ray* rays, size_t n_rays, float max_t, const float* max_t_per_ray,
                 vsb_hit* hits, int mem, int traverse, void* cuda_stream);

/* vsb_closest_hit over ONE scene and a DEVICE ray list, for callers that cannot allocate (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same kernels, same hit records as vsb_closest_hit(..., VSB_MEM_DEVICE,
 * VSB_TRAVERSE_WIDE, stream); but every byte of scratch (scene table, work counter, the coherence sort's keys, order
 * and radix-sort temporaries) lives in one file-scope __device__ array of VSB_STATIC_SCRATCH_BYTES, so the call
 * performs no cudaMalloc / cudaFree / allocating library call.  Limits and caveats:
 *   - n_rays <= VSB_STATIC_MAX_RAYS, else VSB_ERR_TOO_LARGE ("size beyond the fixed maximum");
 *   - every GPU that loads the library holds the array (24 MiB + 8 KiB) until the process exits;
 *   - all calls share the array: calls that overlap on the GPU race on it.  Calls of one context are ordered by the
 *     library; several host threads must serialise their calls. */
#define VSB_STATIC_MAX_RAYS      (1u << 20)
#define VSB_STATIC_SCRATCH_BYTES (8192u + 16u * VSB_STATIC_MAX_RAYS + (8u << 20))
int  vsb_closest_hit_static(const vsb_scene* scene, const vsb_ray* rays, size_t n_rays, float max_t,
                            vsb_hit* hits, void* cuda_stream);

/* multi_hit<N>(r, begin, end) over a ray list (detail/traverse_linear.inl:325-380 ->
 * intersect<MultiHit, N>(), detail/bvh/intersect.inl:35-134; record handling detail/multi_hit.h):
 * hits[i * max_hits + k] = the k-th closest intersection of ray i, sorted by t, equal t in visit
 * order, misses (hit = 0, t = FLT_MAX) behind; max_hits = 1..16 (the reference's template argument,
 * 16 in src/examples/multi_hit/main.cpp:186).  VSB_TRAVERSE_EXACT: records bit-identical to the
 * reference's; VSB_TRAVERSE_WIDE: same records except order / survivor among exact t-ties. */
int  vsb_multi_hit(const vsb_scene* const* scenes, int n_scenes,
                   const vsb_ray* rays, size_t n_rays, int max_hits,
                   vsb_hit* hits, int mem, int traverse, void* cuda_stream);

/* ---- frame: replaces cuda_sched<R>::frame(kernel, sched_params, frame_num)
 * (detail/cuda_sched.h:24-40, detail/cuda_sched.inl:307-320) for the built-in kernels
 * simple / whitted / pathtracing (detail/*.inl) and the AO kernel of
 * src/examples/ao/main.cpp:183-246. ---- */
int  vsb_camera_basis(vsb_frame_desc* d);        /* fills cam_eye/u/v/w from p.eye/center/up/fovy/aspect */
int  vsb_frame(const vsb_scene* const* scenes, int n_scenes, const vsb_frame_desc* d, void* cuda_stream);
int  vsb_tiles_of_rank(int width, int height, int rank, int nranks, int* n_tiles);
/* scatter rank-major gathered tile buffers (nranks * tiles_per_rank_padded * 4096 px) into a
 * row-major RGBA32F image: the de-interleave step after the NCCL all-gather. device pointers. */
int  vsb_deinterleave_tiles(const void* gathered, int width, int height, int nranks,
                            int tiles_per_rank_padded, void* image, void* cuda_stream);

/* ---- device memory helpers for host adaptors that own device render targets without linking the
 * CUDA runtime themselves (gpu_buffer_rt, gpu_buffer_rt.h:16-48).  dst_mem / src_mem: VSB_MEM_* ---- */
int  vsb_device_alloc(size_t bytes, void** out);
int  vsb_device_free(void* p);
int  vsb_device_copy(void* dst, const void* src, size_t bytes, int dst_mem, int src_mem);

/* ---- sharing a device buffer between the per-GPU processes of one node (CUDA IPC over
 * NVLink/NVSwitch peer memory): export on the owner, open on the peers; handle = 64 bytes ---- */
int  vsb_ipc_export(const void* dev_ptr, void* handle64);
int  vsb_ipc_open(const void* handle64, void** out);
int  vsb_ipc_close(void* p);

/* ---- timing helper for the harness: average device time (ms) of the last query/frame
 * kernels, measured with CUDA events on the stream they were launched on ---- */
int  vsb_last_kernel_ms(float* ms);
int  vsb_launch_count(uint64_t* n);              /* number of kernels this library launched */

#ifdef __cplusplus
}
#endif

#endif /* VSNRAY_B200_H */
