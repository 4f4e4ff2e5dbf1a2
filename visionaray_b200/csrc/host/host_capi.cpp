// host_capi.cpp -- extern "C" wrappers around the host C++ layer (scene generators, binned-SAH
// builder) for the Python harness (tests/, bench.py).  Built into libvsnray_b200_host.so.
// Pure host code: no CUDA, no oracle.

#include <cstdint>
#include <cstring>
#include <vector>

#include "visionaray_b200/bvh.h"
#include "visionaray_b200/obj_loader.h"
#include "scenes.h"
#include "../shade.cuh"   // host instantiation of primary_ray() (same single-rounding arithmetic as the kernels)

using namespace visionaray_b200;

extern "C"
{

// kind: 1 uv_sphere(a=slices, b=stacks), 2 displaced_grid(a=n, seed), 3 cube_field(a=n_cubes, seed),
//       5 soup(a=n, seed, s=fparam)
// returns the number of triangles; writes up to cap triangles (64 B each) when out != NULL
size_t vsbh_gen_triangles(int kind, long a, long b, uint64_t seed, float fparam, void* out, size_t cap)
{
    std::vector<scenes::triangle> tris;
    switch (kind)
    {
    case 1: scenes::uv_sphere(tris, int(a), int(b), fparam > 0.0f ? fparam : 1.0f); break;
    case 2: scenes::displaced_grid(tris, int(a), seed, fparam > 0.0f ? fparam : 0.15f); break;
    case 3: scenes::cube_field(tris, int(a), seed); break;
    case 5: scenes::soup(tris, size_t(a), seed, fparam > 0.0f ? fparam : 0.004f); break;
    default: return 0;
    }
    if (out && cap >= tris.size()) std::memcpy(out, tris.data(), tris.size() * sizeof(scenes::triangle));
    return tris.size();
}

size_t vsbh_gen_spheres(size_t n, uint64_t seed, float r_lo, float r_hi, void* out)
{
    std::vector<scenes::sphere> s;
    scenes::sphere_cloud(s, n, seed, r_lo, r_hi);
    std::memcpy(out, s.data(), n * sizeof(scenes::sphere));
    return n;
}

void vsbh_gen_random_rays(void* out, size_t n, uint64_t seed, float lo, float hi)
{
    scenes::random_rays(reinterpret_cast<ray*>(out), n, seed, lo, hi);
}

void vsbh_face_normals(const void* tris, size_t n, void* normals16)
{
    auto t = reinterpret_cast<scenes::triangle const*>(tris);
    auto o = reinterpret_cast<vec3*>(normals16);
    for (size_t i = 0; i < n; ++i) o[i] = normalize(cross(t[i].e1, t[i].e2));
}

// build<index_bvh<P>>(); prim_kind 0 triangles, 1 spheres.  nodes_out needs 2n-1 entries,
// idx_out n entries.
int vsbh_build_bvh_ex(int prim_kind, const void* prims, size_t n, int spatial_splits, void* nodes_out, size_t nodes_cap,
                      size_t* n_nodes, uint32_t* idx_out, size_t idx_cap, size_t* n_idx);

// spot_light<T>::set_spot_cutoff keeps cos(cutoff) (detail/spot_light.inl:143-149): the host's cosf, as the adaptor's setter
float vsbh_spot_cos_cutoff(float cutoff_radians) { return std::cos(cutoff_radians); }

int vsbh_build_bvh(int prim_kind, const void* prims, size_t n, void* nodes_out, size_t nodes_cap,
                   size_t* n_nodes, uint32_t* idx_out, size_t idx_cap, size_t* n_idx)
{
    return vsbh_build_bvh_ex(prim_kind, prims, n, 0, nodes_out, nodes_cap, n_nodes, idx_out, idx_cap, n_idx);
}

// ... with the reference's third argument: build<index_bvh<P>>(first, last, enable_spatial_splits).
// A split tree has more indices than primitives (and up to 2 * indices - 1 nodes): size the outputs for it.
int vsbh_build_bvh_ex(int prim_kind, const void* prims, size_t n, int spatial_splits, void* nodes_out, size_t nodes_cap,
                      size_t* n_nodes, uint32_t* idx_out, size_t idx_cap, size_t* n_idx)
{
    try
    {
        if (prim_kind == 0)
        {
            auto t = build<index_bvh<basic_triangle<3, float>>>(reinterpret_cast<basic_triangle<3, float> const*>(prims), n, spatial_splits != 0);
            *n_nodes = t.num_nodes(); *n_idx = t.num_indices();
            if (*n_nodes > nodes_cap || *n_idx > idx_cap) return 1;
            std::memcpy(nodes_out, t.nodes().data(), *n_nodes * sizeof(bvh_node));
            std::memcpy(idx_out, t.indices().data(), *n_idx * sizeof(uint32_t));
        }
        else if (prim_kind == 2)
        {
            auto t = build<index_bvh<generic_triangle_sphere>>(reinterpret_cast<generic_triangle_sphere const*>(prims), n, spatial_splits != 0);
            *n_nodes = t.num_nodes(); *n_idx = t.num_indices();
            if (*n_nodes > nodes_cap || *n_idx > idx_cap) return 1;
            std::memcpy(nodes_out, t.nodes().data(), *n_nodes * sizeof(bvh_node));
            std::memcpy(idx_out, t.indices().data(), *n_idx * sizeof(uint32_t));
        }
        else
        {
            auto t = build<index_bvh<basic_sphere<float>>>(reinterpret_cast<basic_sphere<float> const*>(prims), n, spatial_splits != 0);
            *n_nodes = t.num_nodes(); *n_idx = t.num_indices();
            if (*n_nodes > nodes_cap || *n_idx > idx_cap) return 1;
            std::memcpy(nodes_out, t.nodes().data(), *n_nodes * sizeof(bvh_node));
            std::memcpy(idx_out, t.indices().data(), *n_idx * sizeof(uint32_t));
        }
    }
    catch (...)
    {
        return 2;
    }
    return 0;
}

// Primary rays of a width x height pinhole frame, uniform pixel sampler, in pixel order
// (ray i = pixel (i % width, i / width)); basis = eye, u, v, w as vsb_camera_basis() fills them
// (4 floats each).  Same arithmetic as the frame kernels' ray generation
// (detail/sched_common.h:130-150), so a ray list made here equals the rays a frame traces.
void vsbh_primary_rays(const float* eye, const float* cu, const float* cv, const float* cw, int width, int height, void* out)
{
    vsb::FrameCtx f;
    std::memset(&f, 0, sizeof(f));
    f.p.width = width; f.p.height = height;
    for (int i = 0; i < 3; ++i) { f.eye[i] = eye[i]; f.cu[i] = cu[i]; f.cv[i] = cv[i]; f.cw[i] = cw[i]; }
    vsb::Ray32* r = reinterpret_cast<vsb::Ray32*>(out);
    for (int y = 0; y < height; ++y)
        for (int x = 0; x < width; ++x)
        {
            vsb::V3 o, d;
            vsb::primary_ray(f, (float)x, (float)y, o, d);
            vsb::Ray32& q = r[(size_t)y * width + x];
            q.ox = o.x; q.oy = o.y; q.oz = o.z; q.pad0 = 0.0f; q.dx = d.x; q.dy = d.y; q.dz = d.z; q.pad1 = 0.0f;
        }
}

// ---- load_obj (include/visionaray_b200/obj_loader.h) for the Python harness -------------------------------------
void* vsbh_obj_load(const char* path)
{
    model* m = new model;
    try { load_obj(path, *m); }
    catch (std::exception const&) { delete m; return nullptr; }
    return m;
}
void vsbh_obj_free(void* h) { delete static_cast<model*>(h); }
// out: primitives, shading normals, geometric normals, tex coords, materials, textures, warnings
void vsbh_obj_counts(void* h, size_t* out7)
{
    model const& m = *static_cast<model*>(h);
    out7[0] = m.primitives.size(); out7[1] = m.shading_normals.size(); out7[2] = m.geometric_normals.size();
    out7[3] = m.tex_coords.size(); out7[4] = m.materials.size(); out7[5] = m.textures.size(); out7[6] = m.warnings.size();
}
// any pointer may be NULL; normals are 16-byte vec3, tex coords 8-byte vec2, materials 64-byte vsb_material, bbox 2 x vec3 (32 B)
void vsbh_obj_copy(void* h, void* prims64, void* shading16, void* geometric16, void* tex8, void* materials64, void* bbox32)
{
    model const& m = *static_cast<model*>(h);
    static_assert(sizeof(model::triangle_type) == 64 && sizeof(vec3) == 16 && sizeof(vec2) == 8, "layouts");
    if (prims64) std::memcpy(prims64, m.primitives.data(), m.primitives.size() * 64);
    if (shading16) std::memcpy(shading16, m.shading_normals.data(), m.shading_normals.size() * 16);
    if (geometric16) std::memcpy(geometric16, m.geometric_normals.data(), m.geometric_normals.size() * 16);
    if (tex8) std::memcpy(tex8, m.tex_coords.data(), m.tex_coords.size() * 8);
    if (materials64)
        for (size_t i = 0; i < m.materials.size(); ++i) { vsb_material f = m.materials[i].flatten(); std::memcpy(static_cast<char*>(materials64) + 64 * i, &f, 64); }
    if (bbox32) { std::memcpy(bbox32, &m.bbox.min, 16); std::memcpy(static_cast<char*>(bbox32) + 16, &m.bbox.max, 16); }
}
// texture i: size, modes, texels (RGBA8, owned by the model); returns 0 for an empty ref
int vsbh_obj_texture(void* h, size_t i, int* width, int* height, int* filter, int* address2, const void** texels)
{
    model const& m = *static_cast<model*>(h);
    if (i >= m.textures.size()) return 0;
    texture_ref2d const& t = m.textures[i];
    *width = int(t.width()); *height = int(t.height()); *filter = int(t.get_filter_mode());
    address2[0] = int(t.get_address_mode(0)); address2[1] = int(t.get_address_mode(1));
    *texels = t.data();
    return t.data() != nullptr && t.width() > 0 && t.height() > 0;
}
size_t vsbh_obj_warning(void* h, size_t i, char* out, size_t cap)
{
    model const& m = *static_cast<model*>(h);
    if (i >= m.warnings.size() || cap == 0) return 0;
    size_t n = m.warnings[i].size() < cap - 1 ? m.warnings[i].size() : cap - 1;
    std::memcpy(out, m.warnings[i].data(), n); out[n] = 0;
    return n;
}

} // extern "C"
