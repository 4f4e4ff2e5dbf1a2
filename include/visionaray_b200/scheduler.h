// visionaray_b200/scheduler.h -- host C++ adaptor of the drop-in boundary: the reference's
// source-level contracts for the ray-query path, with the same names,
// argument meaning and error behaviour, terminating in the C ABI of libvsnray_b200.so
// (include/vsnray_b200.h).
//
//   reference (include/visionaray/...)                         here
//   cuda_index_bvh<P>(index_bvh<P> const&)  bvh.h:344-350      b200_index_bvh<P>(index_bvh<P> const&)
//   cuda_bvh<P>                             bvh.h:443-448      b200_bvh<P>
//   tree.ref()                              bvh.h:364-376      b200_bvh_ref (POD handle)
//   closest_hit / any_hit(ray, b, e[,max_t]) detail/traverse_linear.inl:186-318   same names (+ ray-list overloads)
//   hit_record<ray, primitive<unsigned>>    math/intersect.h:88-115, detail/bvh/hit_record.h:30-46   hit_record
//   pinhole_camera                          pinhole_camera.h   pinhole_camera
//   plastic<T> / matte<T> / point_light<T>  material.h, point_light.h              same names
//   make_kernel_params(...)                 kernels.h:328-509  same name (normals per face / per vertex, tex coords + textures, colours)
//   emissive / mirror / generic_material     material.h, generic_material.h  same names; texture_ref2d = texture_ref<vector<4, unorm<8>>, 2>
//   simple/whitted/pathtracing::kernel<KP>  detail/*.inl       same names (+ ao::kernel, src/examples/ao)
//   pixel_sampler::*_type                   detail/sched_common.h:29-52            same names
//   make_sched_params(...)                  scheduler.h:164-242                    same name
//   cuda_sched<R>::frame(K, SP, frame_num)  detail/cuda_sched.h:24-40              b200_sched<R>::frame
//   gpu_buffer_rt / simple_buffer_rt        gpu_buffer_rt.h, simple_buffer_rt.h    same names
//
// Errors: the reference's path reports none (asserts only); setup errors are std::runtime_error
// there.  Here every failing C-ABI call throws std::runtime_error with vsb_strerror().
// User-defined kernel functors cannot cross the C ABI; they stay on the reference's own template
// path (documented out of scope).
#pragma once
#ifndef VISIONARAY_B200_SCHEDULER_H
#define VISIONARAY_B200_SCHEDULER_H 1

#include <cstring>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <array>
#include <vector>

#include "../vsnray_b200.h"
#include "bvh.h"
#include "math.h"

namespace visionaray_b200
{

namespace detail
{
inline void check(int rc, char const* what)
{
    if (rc != VSB_OK) throw std::runtime_error(std::string(what) + ": " + vsb_strerror(rc));
}
inline void ensure_init()
{
    static bool done = false;
    if (!done) { check(vsb_init(0), "vsb_init"); done = true; }
}
template <typename P> struct prim_kind;
template <> struct prim_kind<basic_triangle<3, float>> { static const int value = VSB_PRIM_TRIANGLE; };
template <> struct prim_kind<basic_sphere<float>> { static const int value = VSB_PRIM_SPHERE; };
template <> struct prim_kind<generic_triangle_sphere> { static const int value = VSB_PRIM_GENERIC; };
} // namespace detail

//-------------------------------------------------------------------------------------------------
// Device BVH: upload + GPU re-encode (cuda_index_bvh<P> / cuda_bvh<P>, bvh.h:344-350,443-448)
//

struct b200_bvh_ref          // what .ref() hands to kernels; trivially copyable
{
    vsb_scene const* scene;
};

template <typename P, bool Indexed>
class b200_bvh_base
{
public:
    using primitive_type = P;
    using bvh_ref = b200_bvh_ref;

    b200_bvh_base() = default;
    b200_bvh_base(b200_bvh_base const&) = delete;
    b200_bvh_base& operator=(b200_bvh_base const&) = delete;
    b200_bvh_base(b200_bvh_base&& o) noexcept : scene_(o.scene_), n_prims_(o.n_prims_), n_nodes_(o.n_nodes_) { o.scene_ = nullptr; }
    b200_bvh_base& operator=(b200_bvh_base&& o) noexcept
    {
        if (this != &o) { reset(); scene_ = o.scene_; n_prims_ = o.n_prims_; n_nodes_ = o.n_nodes_; o.scene_ = nullptr; }
        return *this;
    }
    ~b200_bvh_base() { reset(); }

    bvh_ref ref() const { return bvh_ref{ scene_ }; }
    size_t num_primitives() const { return n_prims_; }
    size_t num_nodes() const { return n_nodes_; }

    // per-face normals kept with the device tree (what viewer.cpp:507-521 keeps in a
    // thrust::device_vector and hands to make_kernel_params)
    void set_normals(vec3 const* normals, size_t count)
    {
        detail::check(vsb_scene_set_normals(scene_, reinterpret_cast<float const*>(normals), count, VSB_MEM_HOST), "vsb_scene_set_normals");
    }

    // normals with their binding, colours, tex coords and textures kept with the device tree (deep copy);
    // frames whose kernel_params pass nullptr normals then use them
    template <typename SP /* surface_params */>
    void set_surface(SP const& sp)
    {
        std::vector<vsb_texture> texs;
        vsb_surface_desc d = sp.flatten(texs);
        detail::check(vsb_scene_set_surface(scene_, sp.present ? &d : nullptr, VSB_MEM_HOST), "vsb_scene_set_surface");
    }

protected:
    void create(P const* prims, size_t np, bvh_node const* nodes, size_t nn, unsigned const* idx, size_t ni)
    {
        detail::ensure_init();
        reset();
        detail::check(vsb_scene_create(detail::prim_kind<P>::value, prims, np, reinterpret_cast<vsb_bvh_node const*>(nodes), nn,
                                       idx, ni, &scene_), "vsb_scene_create");
        n_prims_ = np; n_nodes_ = nn;
    }
    void reset() { if (scene_) { vsb_scene_destroy(scene_); scene_ = nullptr; } }

    vsb_scene* scene_ = nullptr;
    size_t n_prims_ = 0, n_nodes_ = 0;
};

template <typename P>
class b200_index_bvh : public b200_bvh_base<P, true>
{
public:
    b200_index_bvh() = default;
    explicit b200_index_bvh(index_bvh<P> const& rhs)
    {
        this->create(rhs.primitives().data(), rhs.num_primitives(), rhs.nodes().data(), rhs.num_nodes(),
                     rhs.indices().data(), rhs.num_indices());
    }
    // raw arrays in the reference's layouts (e.g. straight out of the reference's own index_bvh)
    b200_index_bvh(P const* prims, size_t np, bvh_node const* nodes, size_t nn, unsigned const* idx, size_t ni)
    {
        this->create(prims, np, nodes, nn, idx, ni);
    }

    // The tree BUILT ON THE GPU from the primitives alone (vsb_scene_build): what
    // build<index_bvh<P>>(first, last) + cuda_index_bvh<P>(host_bvh) do in the reference
    // (bvh.h:455,344-350; detail/bvh/build.inl, sah.h), without the single-threaded host build.
    // Spelled like the reference's builder: build<b200_index_bvh<P>>(prims, n).
    static b200_index_bvh build_on_device(P const* prims, size_t np)
    {
        detail::ensure_init();
        b200_index_bvh out;
        detail::check(vsb_scene_build(detail::prim_kind<P>::value, prims, np, VSB_MEM_HOST, &out.scene_), "vsb_scene_build");
        uint64_t st[6] = { 0, 0, 0, 0, 0, 0 };
        detail::check(vsb_scene_stats(out.scene_, st, 6), "vsb_scene_stats");
        out.n_prims_ = np; out.n_nodes_ = size_t(st[5]);
        return out;
    }

    // the device tree back as a host index_bvh<P> in the reference's layout (nodes(), indices());
    // `prims` are the caller's primitives (the device copy is not read back)
    index_bvh<P> download(P const* prims) const
    {
        index_bvh<P> host(prims, this->n_prims_);
        host.nodes().resize(this->n_nodes_);
        host.indices().resize(this->n_prims_);
        detail::check(vsb_scene_download_bvh(this->scene_, reinterpret_cast<vsb_bvh_node*>(host.nodes().data()), host.nodes().size(),
                                             host.indices().data(), host.indices().size()), "vsb_scene_download_bvh");
        return host;
    }
};

template <typename T> struct is_b200_index_bvh : std::false_type {};
template <typename P> struct is_b200_index_bvh<b200_index_bvh<P>> : std::true_type {};

// build<b200_index_bvh<P>>(prims, n): same spelling as build<index_bvh<P>>(prims, n) (bvh.h:455)
template <typename Tree, typename P>
typename std::enable_if<is_b200_index_bvh<Tree>::value, Tree>::type build(P const* prims, size_t n)
{
    return Tree::build_on_device(prims, n);
}

template <typename P>
class b200_bvh : public b200_bvh_base<P, false>
{
public:
    b200_bvh() = default;
    explicit b200_bvh(bvh<P> const& rhs)
    {
        this->create(rhs.primitives().data(), rhs.num_primitives(), rhs.nodes().data(), rhs.num_nodes(), nullptr, 0);
    }
};

//-------------------------------------------------------------------------------------------------
// Ray queries (detail/traverse_linear.inl:186-318)
//

// hit_record_bvh<ray, BVH, hit_record<ray, primitive<unsigned>>>; same fields, same miss values
struct hit_record
{
    int      hit;
    unsigned prim_id;
    unsigned geom_id;
    float    t;
    float    u;
    float    v;
    unsigned primitive_list_index;
    unsigned pad_;
    explicit operator bool() const { return hit != 0; }
};
static_assert(sizeof(hit_record) == sizeof(vsb_hit), "hit record layout");

namespace detail
{
inline void query(bool any, b200_bvh_ref const* begin, b200_bvh_ref const* end, ray const* rays, size_t n,
                  float max_t, float const* max_t_per_ray, hit_record* hits, int traverse)
{
    std::vector<vsb_scene const*> s;
    for (auto it = begin; it != end; ++it) s.push_back(it->scene);
    auto fn = any ? vsb_any_hit : vsb_closest_hit;
    check(fn(s.data(), int(s.size()), reinterpret_cast<vsb_ray const*>(rays), n, max_t, max_t_per_ray,
             reinterpret_cast<vsb_hit*>(hits), VSB_MEM_HOST, traverse, nullptr), any ? "vsb_any_hit" : "vsb_closest_hit");
}
} // namespace detail

// single ray, as in the reference (one launch per call: for real work use the ray-list overloads)
inline hit_record closest_hit(ray const& r, b200_bvh_ref const* begin, b200_bvh_ref const* end)
{
    hit_record h;
    detail::query(false, begin, end, &r, 1, FLT_MAX, nullptr, &h, VSB_TRAVERSE_WIDE);
    return h;
}
inline hit_record any_hit(ray const& r, b200_bvh_ref const* begin, b200_bvh_ref const* end, float max_t = FLT_MAX)
{
    hit_record h;
    detail::query(true, begin, end, &r, 1, max_t, nullptr, &h, VSB_TRAVERSE_WIDE);
    return h;
}

// ray lists (host memory): hits[i] = closest_hit(rays[i], begin, end)
inline void closest_hit(ray const* rays, size_t n, b200_bvh_ref const* begin, b200_bvh_ref const* end, hit_record* hits,
                        bool reference_order = false)
{
    detail::query(false, begin, end, rays, n, FLT_MAX, nullptr, hits, reference_order ? VSB_TRAVERSE_EXACT : VSB_TRAVERSE_WIDE);
}
inline void any_hit(ray const* rays, size_t n, b200_bvh_ref const* begin, b200_bvh_ref const* end, float max_t, hit_record* hits,
                    bool reference_order = false)
{
    detail::query(true, begin, end, rays, n, max_t, nullptr, hits, reference_order ? VSB_TRAVERSE_EXACT : VSB_TRAVERSE_WIDE);
}

// multi_hit<N>(r, begin, end) (detail/traverse_linear.inl:325-380): the N closest intersections sorted
// by t, misses (hit = 0, t = FLT_MAX) behind; N = 1..16 (16 in src/examples/multi_hit/main.cpp:186)
template <size_t N = 16>
inline std::array<hit_record, N> multi_hit(ray const& r, b200_bvh_ref const* begin, b200_bvh_ref const* end)
{
    static_assert(N >= 1 && N <= 16, "multi_hit<N>: N = 1..16");
    std::array<hit_record, N> h;
    std::vector<vsb_scene const*> scenes;
    for (auto it = begin; it != end; ++it) scenes.push_back(it->scene);
    detail::check(vsb_multi_hit(scenes.data(), int(scenes.size()), reinterpret_cast<vsb_ray const*>(&r), 1, int(N),
                                reinterpret_cast<vsb_hit*>(h.data()), VSB_MEM_HOST, VSB_TRAVERSE_WIDE, nullptr), "vsb_multi_hit");
    return h;
}
// ray list (host memory): hits[i * N + k] = multi_hit<N>(rays[i], begin, end)[k]
template <size_t N = 16>
inline void multi_hit(ray const* rays, size_t n, b200_bvh_ref const* begin, b200_bvh_ref const* end, hit_record* hits,
                      bool reference_order = false)
{
    static_assert(N >= 1 && N <= 16, "multi_hit<N>: N = 1..16");
    std::vector<vsb_scene const*> scenes;
    for (auto it = begin; it != end; ++it) scenes.push_back(it->scene);
    detail::check(vsb_multi_hit(scenes.data(), int(scenes.size()), reinterpret_cast<vsb_ray const*>(rays), n, int(N),
                                reinterpret_cast<vsb_hit*>(hits), VSB_MEM_HOST,
                                reference_order ? VSB_TRAVERSE_EXACT : VSB_TRAVERSE_WIDE, nullptr), "vsb_multi_hit");
}

//-------------------------------------------------------------------------------------------------
// Camera (camera.h:50-94), materials (material.h), point light (point_light.h)
//

class pinhole_camera
{
public:
    // view / projection matrices as the reference camera keeps them (detail/camera.inl:10-65)
    void look_at(vec3 const& eye, vec3 const& center, vec3 const& up = vec3(0.0f, 1.0f, 0.0f))
    {
        eye_ = eye; center_ = center; up_ = up;
        vec3 f = normalize(eye - center);
        vec3 s = normalize(cross(up, f));
        vec3 u = cross(f, s);
        view_ = mat4(s.x, u.x, f.x, 0.0f, s.y, u.y, f.y, 0.0f, s.z, u.z, f.z, 0.0f, -dot(eye, s), -dot(eye, u), -dot(eye, f), 1.0f);
    }
    void perspective(float fovy, float aspect, float z_near, float z_far)
    {
        fovy_ = fovy; aspect_ = aspect; z_near_ = z_near; z_far_ = z_far;
        float f = 1.0f / std::tan(fovy * 0.5f);
        for (int i = 0; i < 16; ++i) proj_.m[i] = 0.0f;
        proj_(0, 0) = f / aspect;
        proj_(1, 1) = f;
        proj_(2, 2) = (z_far + z_near) / (z_near - z_far);
        proj_(2, 3) = (2.0f * z_far * z_near) / (z_near - z_far);
        proj_(3, 2) = -1.0f;
    }
    mat4 const& get_view_matrix() const { return view_; }
    mat4 const& get_proj_matrix() const { return proj_; }
    void set_viewport(recti const& viewport) { viewport_ = viewport; }
    void set_viewport(int left, int bottom, int width, int height) { viewport_ = recti(left, bottom, width, height); }
    recti get_viewport() const { return viewport_; }
    float fovy() const { return fovy_; }
    float aspect() const { return aspect_; }
    float z_near() const { return z_near_; }
    float z_far() const { return z_far_; }
    vec3 const& eye() const { return eye_; }
    vec3 const& center() const { return center_; }
    vec3 const& up() const { return up_; }

private:
    vec3 eye_ = vec3(0.0f, 0.0f, 1.0f), center_ = vec3(0.0f), up_ = vec3(0.0f, 1.0f, 0.0f);
    float fovy_ = 0.78539816f, aspect_ = 1.0f, z_near_ = 0.001f, z_far_ = 1000.0f;
    recti viewport_ = recti(0, 0, 0, 0);
    mat4 view_ = mat4::identity(), proj_ = mat4::identity();
};
using camera = pinhole_camera;   // the reference's class name (camera.h)

// plastic<T> / matte<T>: same getters/setters; flattened to vsb_material at frame() time
template <typename T = float>
class plastic
{
public:
    void set_ca(vec3 const& c) { ca_ = c; } vec3 get_ca() const { return ca_; }
    void set_cd(vec3 const& c) { cd_ = c; } vec3 get_cd() const { return cd_; }
    void set_cs(vec3 const& c) { cs_ = c; } vec3 get_cs() const { return cs_; }
    void set_ka(T k) { ka_ = k; } T get_ka() const { return ka_; }
    void set_kd(T k) { kd_ = k; } T get_kd() const { return kd_; }
    void set_ks(T k) { ks_ = k; } T get_ks() const { return ks_; }
    void set_specular_exp(T e) { exp_ = e; } T get_specular_exp() const { return exp_; }
    vsb_material flatten() const
    {
        vsb_material m; std::memset(&m, 0, sizeof(m));
        for (int i = 0; i < 3; ++i) { m.ca[i] = ca_[i]; m.cd[i] = cd_[i]; m.cs[i] = cs_[i]; }
        m.ka = ka_; m.kd = kd_; m.ks = ks_; m.spec_exp = exp_; m.kind = VSB_MAT_PLASTIC;
        return m;
    }
private:
    vec3 ca_ = vec3(0.0f), cd_ = vec3(0.0f), cs_ = vec3(0.0f);
    T ka_ = T(0), kd_ = T(0), ks_ = T(0), exp_ = T(0);
};

template <typename T = float>
class matte
{
public:
    void set_ca(vec3 const& c) { ca_ = c; } vec3 get_ca() const { return ca_; }
    void set_cd(vec3 const& c) { cd_ = c; } vec3 get_cd() const { return cd_; }
    void set_ka(T k) { ka_ = k; } T get_ka() const { return ka_; }
    void set_kd(T k) { kd_ = k; } T get_kd() const { return kd_; }
    vsb_material flatten() const
    {
        vsb_material m; std::memset(&m, 0, sizeof(m));
        for (int i = 0; i < 3; ++i) { m.ca[i] = ca_[i]; m.cd[i] = cd_[i]; }
        m.ka = ka_; m.kd = kd_; m.kind = VSB_MAT_MATTE;
        return m;
    }
private:
    vec3 ca_ = vec3(0.0f), cd_ = vec3(0.0f);
    T ka_ = T(0), kd_ = T(0);
};

// emissive<T> / mirror<T> (material.h, detail/material/{emissive,mirror}.inl)
template <typename T = float>
class emissive
{
public:
    void set_ce(vec3 const& c) { ce_ = c; } vec3 get_ce() const { return ce_; }
    void set_ls(T ls) { ls_ = ls; } T get_ls() const { return ls_; }
    vsb_material flatten() const
    {
        vsb_material m; std::memset(&m, 0, sizeof(m));
        for (int i = 0; i < 3; ++i) m.cd[i] = ce_[i];
        m.kd = ls_; m.kind = VSB_MAT_EMISSIVE;
        return m;
    }
private:
    vec3 ce_ = vec3(0.0f);
    T ls_ = T(0);
};

template <typename T = float>
class mirror
{
public:
    void set_cr(vec3 const& c) { cr_ = c; } vec3 get_cr() const { return cr_; }
    void set_kr(T k) { kr_ = k; } T get_kr() const { return kr_; }
    void set_ior(vec3 const& c) { ior_ = c; } void set_ior(T v) { ior_ = vec3(v); } vec3 get_ior() const { return ior_; }
    void set_absorption(vec3 const& c) { abs_ = c; } void set_absorption(T v) { abs_ = vec3(v); } vec3 get_absorption() const { return abs_; }
    vsb_material flatten() const
    {
        vsb_material m; std::memset(&m, 0, sizeof(m));
        for (int i = 0; i < 3; ++i) { m.cs[i] = cr_[i]; m.ca[i] = ior_[i]; m.cd[i] = abs_[i]; }
        m.ks = kr_; m.kind = VSB_MAT_MIRROR;
        return m;
    }
private:
    vec3 cr_ = vec3(0.0f), ior_ = vec3(0.0f), abs_ = vec3(0.0f);
    T kr_ = T(0);
};

// generic_material<emissive<T>, matte<T>, mirror<T>, plastic<T>> (generic_material.h:22-70): a variant of
// the four; the flattened record carries the alternative in `kind`
template <typename ...Ts>
class generic_material
{
public:
    generic_material() { std::memset(&m_, 0, sizeof(m_)); m_.kind = VSB_MAT_MATTE; }
    template <typename M, typename = decltype(std::declval<M const&>().flatten())>
    generic_material(M const& material) : m_(material.flatten()) {}
    bool is_emissive() const { return m_.kind == VSB_MAT_EMISSIVE; }
    vsb_material flatten() const { return m_; }
private:
    vsb_material m_;
};

// texture_ref<vector<4, unorm<8>>, 2> (texture/detail/texture2d.h, texture_common.h:36-92,172-215): a
// non-owning view of RGBA8 texels with filter and address modes
enum tex_address_mode { Wrap = 0, Mirror, Clamp };
enum tex_filter_mode { Nearest = 0, Linear };
struct unorm8x4 { unsigned char x, y, z, w; };

class texture_ref2d
{
public:
    texture_ref2d() = default;
    texture_ref2d(size_t w, size_t h) : width_(w), height_(h) {}
    void reset(unorm8x4 const* data) { data_ = data; }
    unorm8x4 const* data() const { return data_; }
    size_t width() const { return width_; }
    size_t height() const { return height_; }
    void set_address_mode(size_t index, tex_address_mode mode) { address_[index] = mode; }
    void set_address_mode(tex_address_mode mode) { address_[0] = address_[1] = mode; }
    tex_address_mode get_address_mode(size_t index) const { return address_[index]; }
    void set_filter_mode(tex_filter_mode mode) { filter_ = mode; }
    tex_filter_mode get_filter_mode() const { return filter_; }
    vsb_texture flatten() const
    {
        vsb_texture t; std::memset(&t, 0, sizeof(t));
        t.data = data_; t.width = int(width_); t.height = int(height_);
        t.filter = int(filter_); t.address[0] = int(address_[0]); t.address[1] = int(address_[1]);
        return t;
    }
private:
    unorm8x4 const* data_ = nullptr;
    size_t width_ = 0, height_ = 0;
    tex_address_mode address_[2] = { Wrap, Wrap };
    tex_filter_mode filter_ = Nearest;
};

// binding tags (tags.h)
struct normals_per_face_binding   { static const int id = VSB_NORMALS_PER_FACE; };
struct normals_per_vertex_binding { static const int id = VSB_NORMALS_PER_VERTEX; };
struct colors_per_face_binding    { static const int id = VSB_COLORS_PER_FACE; };
struct colors_per_vertex_binding  { static const int id = VSB_COLORS_PER_VERTEX; };

// The members of kernel_params beyond prims / materials / lights (kernels.h:194-319): host arrays.
// present == false: per-face normals only (the first two make_kernel_params overloads).
struct surface_params
{
    bool present;
    int normal_binding, color_binding;
    vec3 const* normals;
    vec3 const* colors;
    vec2 const* tex_coords;
    texture_ref2d const* textures; size_t num_textures;

    vsb_surface_desc flatten(std::vector<vsb_texture>& texs) const
    {
        vsb_surface_desc d; std::memset(&d, 0, sizeof(d));
        d.normal_binding = normal_binding; d.color_binding = color_binding;
        d.normals = reinterpret_cast<float const*>(normals);
        d.colors = reinterpret_cast<float const*>(colors);
        d.tex_coords = reinterpret_cast<float const*>(tex_coords);
        texs.clear();
        for (size_t i = 0; i < num_textures; ++i) texs.push_back(textures[i].flatten());
        d.textures = texs.data(); d.n_textures = int(texs.size());
        return d;
    }
};

template <typename T = float>
class point_light
{
public:
    vec3 position() const { return position_; }
    void set_cl(vec3 const& cl) { cl_ = cl; }
    void set_kl(T kl) { kl_ = kl; }
    void set_position(vec3 const& pos) { position_ = pos; }
    void set_constant_attenuation(T att) { att_const_ = att; }
    void set_linear_attenuation(T att) { att_lin_ = att; }
    void set_quadratic_attenuation(T att) { att_quad_ = att; }
    vsb_point_light flatten() const
    {
        vsb_point_light l; std::memset(&l, 0, sizeof(l));
        for (int i = 0; i < 3; ++i) { l.pos[i] = position_[i]; l.cl[i] = cl_[i]; }
        l.kl = kl_; l.att_const = att_const_; l.att_lin = att_lin_; l.att_quad = att_quad_;
        return l;
    }
private:
    vec3 cl_ = vec3(1.0f), position_ = vec3(0.0f);
    T kl_ = T(1), att_const_ = T(1), att_lin_ = T(0), att_quad_ = T(0);
};

// spot_light<T> (spot_light.h, detail/spot_light.inl): set_spot_direction takes a unit vector, set_spot_cutoff an
// angle and keeps its cosine
template <typename T = float>
class spot_light
{
public:
    vec3 position() const { return position_; }
    void set_cl(vec3 const& cl) { cl_ = cl; }
    void set_kl(T kl) { kl_ = kl; }
    void set_position(vec3 const& pos) { position_ = pos; }
    void set_spot_direction(vec3 const& dir) { spot_direction_ = dir; }
    void set_spot_cutoff(T cutoff) { spot_cutoff_ = cutoff; spot_cos_cutoff_ = std::cos(cutoff); }
    void set_spot_exponent(T e) { spot_exponent_ = e; }
    void set_constant_attenuation(T att) { att_const_ = att; }
    void set_linear_attenuation(T att) { att_lin_ = att; }
    void set_quadratic_attenuation(T att) { att_quad_ = att; }
    vec3 spot_direction() const { return spot_direction_; }
    T spot_cutoff() const { return spot_cutoff_; }
    T spot_exponent() const { return spot_exponent_; }
    vsb_point_light flatten() const
    {
        vsb_point_light l; std::memset(&l, 0, sizeof(l));
        for (int i = 0; i < 3; ++i) { l.pos[i] = position_[i]; l.cl[i] = cl_[i]; l.spot_dir[i] = spot_direction_[i]; }
        l.kl = kl_; l.att_const = att_const_; l.att_lin = att_lin_; l.att_quad = att_quad_;
        l.kind = VSB_LIGHT_SPOT; l.spot_exponent = spot_exponent_; l.spot_cos_cutoff = spot_cos_cutoff_;
        return l;
    }
private:
    vec3 cl_ = vec3(1.0f), position_ = vec3(0.0f), spot_direction_ = vec3(0.0f, 0.0f, -1.0f);
    T kl_ = T(1), spot_cutoff_ = T(0), spot_cos_cutoff_ = T(1), spot_exponent_ = T(0);
    T att_const_ = T(1), att_lin_ = T(0), att_quad_ = T(0);
};

//-------------------------------------------------------------------------------------------------
// Kernel parameters and the built-in kernels (kernels.h:94-509, detail/{simple,whitted,pathtracing}.inl)
//

namespace detail
{
template <typename Light>
inline std::vector<vsb_point_light> flatten_lights(Light const* b, Light const* e)
{
    std::vector<vsb_point_light> out;
    for (; b != e; ++b) out.push_back(b->flatten());
    return out;
}
} // namespace detail

template <typename Material>
struct kernel_params
{
    b200_bvh_ref const* prims_begin;
    b200_bvh_ref const* prims_end;
    vec3 const* normals;             // host per-face normals, or nullptr = the normals attached to the device tree
    Material const* materials;  size_t num_materials;     // indexed by geom_id
    std::vector<vsb_point_light> lights;          // point_light<T> / spot_light<T> list, flattened when the params are made
    unsigned num_bounces;
    float epsilon;
    vec4 bg_color;
    vec4 ambient_color;
    // normals per vertex / colours / tex coords + textures (the later make_kernel_params overloads).  Host
    // arrays given here are uploaded by every frame(); attach them once with b200_index_bvh::set_surface()
    // and pass nullptr normals to keep them resident, like the thrust vectors of a cuda_sched user
    surface_params surface;
};

// make_kernel_params(normals_per_face_binding, prims b/e, normals, materials, lights b/e, ...)  kernels.h:371-409
// `materials` is a contiguous container (std::vector<plastic<float>>, ...), as the reference's callers pass.
template <typename Materials, typename Light>
auto make_kernel_params(normals_per_face_binding, b200_bvh_ref const* begin, b200_bvh_ref const* end, vec3 const* normals,
                        Materials const& materials, Light const* lbegin, Light const* lend,
                        unsigned num_bounces = 5, float epsilon = FLT_EPSILON, vec4 const& bg_color = vec4(0.0f),
                        vec4 const& ambient_color = vec4(0.0f))
    -> kernel_params<typename std::decay<decltype(materials[0])>::type>
{
    using M = typename std::decay<decltype(materials[0])>::type;
    return kernel_params<M>{ begin, end, normals, &materials[0], materials.size(), detail::flatten_lights(lbegin, lend), num_bounces, epsilon, bg_color, ambient_color,
                             surface_params{ false, 0, 0, nullptr, nullptr, nullptr, nullptr, 0 } };
}

// make_kernel_params(prims b/e, materials, lights b/e, ...)  kernels.h:328-366 -- no normals: the kernels take
// get_normal(hr, primitive), i.e. normalize(cross(e1, e2)) for triangles and (isect_pos - center) / radius for
// spheres; the form that works for a sphere bvh (simple / whitted kernels)
template <typename Materials, typename Light, typename = decltype(std::declval<Materials const&>()[0].flatten())>
auto make_kernel_params(b200_bvh_ref const* begin, b200_bvh_ref const* end, Materials const& materials,
                        Light const* lbegin, Light const* lend,
                        unsigned num_bounces = 5, float epsilon = FLT_EPSILON, vec4 const& bg_color = vec4(0.0f),
                        vec4 const& ambient_color = vec4(0.0f))
    -> kernel_params<typename std::decay<decltype(materials[0])>::type>
{
    using M = typename std::decay<decltype(materials[0])>::type;
    return kernel_params<M>{ begin, end, nullptr, &materials[0], materials.size(), detail::flatten_lights(lbegin, lend), num_bounces, epsilon, bg_color, ambient_color,
                             surface_params{ true, VSB_NORMALS_FROM_PRIMITIVE, 0, nullptr, nullptr, nullptr, nullptr, 0 } };
}

// make_kernel_params(normals_per_vertex_binding, prims b/e, normals, materials, lights b/e, ...)  kernels.h:371-393
template <typename Materials, typename Light>
auto make_kernel_params(normals_per_vertex_binding, b200_bvh_ref const* begin, b200_bvh_ref const* end, vec3 const* normals,
                        Materials const& materials, Light const* lbegin, Light const* lend,
                        unsigned num_bounces = 5, float epsilon = FLT_EPSILON, vec4 const& bg_color = vec4(0.0f),
                        vec4 const& ambient_color = vec4(0.0f))
    -> kernel_params<typename std::decay<decltype(materials[0])>::type>
{
    using M = typename std::decay<decltype(materials[0])>::type;
    return kernel_params<M>{ begin, end, nullptr, &materials[0], materials.size(), detail::flatten_lights(lbegin, lend), num_bounces, epsilon, bg_color, ambient_color,
                             surface_params{ true, VSB_NORMALS_PER_VERTEX, 0, normals, nullptr, nullptr, nullptr, 0 } };
}

// w/ normals and textures  kernels.h:395-437.  `textures`: contiguous container of texture_ref2d, by geom_id
template <typename NormalBinding, typename Materials, typename Textures, typename Light, int = NormalBinding::id>
auto make_kernel_params(NormalBinding, b200_bvh_ref const* begin, b200_bvh_ref const* end, vec3 const* normals, vec2 const* tex_coords,
                        Materials const& materials, Textures const& textures, Light const* lbegin,
                        Light const* lend, unsigned num_bounces = 5, float epsilon = FLT_EPSILON,
                        vec4 const& bg_color = vec4(0.0f), vec4 const& ambient_color = vec4(0.0f))
    -> kernel_params<typename std::decay<decltype(materials[0])>::type>
{
    using M = typename std::decay<decltype(materials[0])>::type;
    return kernel_params<M>{ begin, end, nullptr, &materials[0], materials.size(), detail::flatten_lights(lbegin, lend), num_bounces, epsilon, bg_color, ambient_color,
                             surface_params{ true, NormalBinding::id, 0, normals, nullptr, tex_coords, &textures[0], textures.size() } };
}

// w/ normals, colors and textures  kernels.h:440-509
template <typename NormalBinding, typename ColorBinding, typename Materials, typename Textures, typename Light, int = NormalBinding::id, int = ColorBinding::id>
auto make_kernel_params(NormalBinding, ColorBinding, b200_bvh_ref const* begin, b200_bvh_ref const* end, vec3 const* normals,
                        vec2 const* tex_coords, Materials const& materials, vec3 const* colors, Textures const& textures,
                        Light const* lbegin, Light const* lend, unsigned num_bounces = 5,
                        float epsilon = FLT_EPSILON, vec4 const& bg_color = vec4(0.0f), vec4 const& ambient_color = vec4(0.0f))
    -> kernel_params<typename std::decay<decltype(materials[0])>::type>
{
    using M = typename std::decay<decltype(materials[0])>::type;
    return kernel_params<M>{ begin, end, nullptr, &materials[0], materials.size(), detail::flatten_lights(lbegin, lend), num_bounces, epsilon, bg_color, ambient_color,
                             surface_params{ true, NormalBinding::id, ColorBinding::id, normals, colors, tex_coords, &textures[0], textures.size() } };
}

namespace simple      { template <typename KP> struct kernel { KP params; static const int id = VSB_KERNEL_SIMPLE; }; }
namespace whitted     { template <typename KP> struct kernel { KP params; static const int id = VSB_KERNEL_WHITTED; }; }
namespace pathtracing { template <typename KP> struct kernel { KP params; static const int id = VSB_KERNEL_PATHTRACING; }; }
// the AO kernel of src/examples/ao/main.cpp:183-246 (a lambda there)
namespace ao
{
template <typename KP> struct kernel { KP params; int num_samples; float radius; static const int id = VSB_KERNEL_AO; };
}

//-------------------------------------------------------------------------------------------------
// Pixel samplers (detail/sched_common.h:29-52) and scheduler params (scheduler.h:23-242)
//

namespace pixel_sampler
{
struct uniform_type        { static const int id = VSB_PS_UNIFORM; };
struct jittered_type       { static const int id = VSB_PS_JITTERED; };
struct jittered_blend_type { static const int id = VSB_PS_JITTERED_BLEND; };
template <unsigned Num> struct ssaa_type
{
    static_assert(Num == 1 || Num == 2 || Num == 4 || Num == 8, "ssaa_type<1|2|4|8>");
    static const int id = Num == 1 ? VSB_PS_UNIFORM : (Num == 2 ? VSB_PS_SSAA2 : (Num == 4 ? VSB_PS_SSAA4 : VSB_PS_SSAA8));
};
}

template <typename RT, typename PxSampler>
struct sched_params
{
    using pixel_sampler_type = PxSampler;
    pinhole_camera cam;
    RT& rt;                    // must outlive frame(), as in the reference
    recti scissor_box;
};

template <typename PxSampler, typename RT>
sched_params<RT, PxSampler> make_sched_params(PxSampler, pinhole_camera const& cam, RT& rt)
{
    return sched_params<RT, PxSampler>{ cam, rt, recti(0, 0, int(rt.width()), int(rt.height())) };
}
template <typename RT>
sched_params<RT, pixel_sampler::uniform_type> make_sched_params(pinhole_camera const& cam, RT& rt)
{
    return make_sched_params(pixel_sampler::uniform_type{}, cam, rt);
}

// sched_params from view / projection matrices (scheduler.h:60-96, make_sched_params :205-242):
// primary rays from the inverse matrices, and the only form that can fill a depth buffer
template <typename RT, typename PxSampler>
struct sched_params_matrix
{
    using pixel_sampler_type = PxSampler;
    mat4 view_matrix, proj_matrix;
    RT& rt;
    recti scissor_box;
};

template <typename PxSampler, typename RT>
sched_params_matrix<RT, PxSampler> make_sched_params(PxSampler, mat4 const& view_matrix, mat4 const& proj_matrix, RT& rt)
{
    return sched_params_matrix<RT, PxSampler>{ view_matrix, proj_matrix, rt, recti(0, 0, int(rt.width()), int(rt.height())) };
}
template <typename RT>
sched_params_matrix<RT, pixel_sampler::uniform_type> make_sched_params(mat4 const& view_matrix, mat4 const& proj_matrix, RT& rt)
{
    return make_sched_params(pixel_sampler::uniform_type{}, view_matrix, proj_matrix, rt);
}

//-------------------------------------------------------------------------------------------------
// Render targets (render_target.h:45-101, simple_buffer_rt.h:21-52, gpu_buffer_rt.h:16-48).
// Colour PF_RGBA32F, PF_RGBA8, PF_RGB32F or PF_RGB8 (pixel_format.h, pixel_traits.h), depth PF_UNSPECIFIED
// (none) or PF_DEPTH32F: buffer[y * width + x], y up, pitch = width.
//

enum pixel_format { PF_UNSPECIFIED = 0, PF_RGBA32F = 1, PF_DEPTH32F = 2, PF_RGBA8 = 3, PF_RGB32F = 4, PF_RGB8 = 5 };

struct unorm8x3 { unsigned char x, y, z; };

// pixel_traits<PF>::type (pixel_traits.h): the element type of a colour buffer of that format and its
// VSB_PF_* id.  vector<3,float> is 16-byte aligned in the reference, hence a PF_RGB32F pixel is a vec3.
template <pixel_format PF> struct pixel_traits;
template <> struct pixel_traits<PF_RGBA32F> { using type = vec4;     static const int id = VSB_PF_RGBA32F; };
template <> struct pixel_traits<PF_RGBA8>   { using type = unorm8x4; static const int id = VSB_PF_RGBA8; };
template <> struct pixel_traits<PF_RGB32F>  { using type = vec3;     static const int id = VSB_PF_RGB32F; };
template <> struct pixel_traits<PF_RGB8>    { using type = unorm8x3; static const int id = VSB_PF_RGB8; };

namespace detail
{
// convert(target, vec4) of pixel_access::store (detail/color_conversion.h:213-247, math/detail/unorm.inl:18-24)
inline unsigned char to_unorm8(float f) { f = f < 0.0f ? 0.0f : (f > 1.0f ? 1.0f : f); return (unsigned char)(unsigned)(f * 255.0); }
template <typename T> T to_pixel(vec4 const& c);
template <> inline vec4 to_pixel<vec4>(vec4 const& c) { return c; }
template <> inline unorm8x4 to_pixel<unorm8x4>(vec4 const& c) { return unorm8x4{ to_unorm8(c.x), to_unorm8(c.y), to_unorm8(c.z), to_unorm8(c.w) }; }
template <> inline vec3 to_pixel<vec3>(vec4 const& c) { return vec3(c.x * c.w, c.y * c.w, c.z * c.w); }
template <> inline unorm8x3 to_pixel<unorm8x3>(vec4 const& c) { return unorm8x3{ to_unorm8(c.x * c.w), to_unorm8(c.y * c.w), to_unorm8(c.z * c.w) }; }
} // namespace detail

template <pixel_format CF, pixel_format DF>
struct render_target_ref
{
    using color_type = typename pixel_traits<CF>::type;
    color_type* color_; float* depth_; int width_, height_;
    color_type* color() const { return color_; }
    float* depth() const { return depth_; }
    int width() const { return width_; }
    int height() const { return height_; }
};

// host memory; frame() returns with the pixels in place (like tiled_sched into simple_buffer_rt)
template <pixel_format CF = PF_RGBA32F, pixel_format DF = PF_UNSPECIFIED>
class simple_buffer_rt
{
    static_assert(DF == PF_UNSPECIFIED || DF == PF_DEPTH32F, "the path writes DEPTH32F depth");
public:
    using color_type = typename pixel_traits<CF>::type;
    using depth_type = float;
    using ref_type = render_target_ref<CF, DF>;
    static const int memory = VSB_MEM_HOST;
    static const int color_format = pixel_traits<CF>::id;

    color_type* color() { return color_.data(); }
    color_type const* color() const { return color_.data(); }
    depth_type* depth() { return DF == PF_DEPTH32F ? depth_.data() : nullptr; }
    depth_type const* depth() const { return DF == PF_DEPTH32F ? depth_.data() : nullptr; }
    ref_type ref() { return ref_type{ color(), depth(), int(width_), int(height_) }; }
    void clear_color_buffer(vec4 const& c = vec4(0.0f)) { color_type q = detail::to_pixel<color_type>(c); for (auto& p : color_) p = q; }
    void clear_depth_buffer(float d = 1.0f) { for (auto& p : depth_) p = d; }
    void begin_frame() {}
    void end_frame() {}
    void resize(size_t w, size_t h)
    {
        width_ = w; height_ = h; color_.assign(w * h, detail::to_pixel<color_type>(vec4(0.0f)));
        if (DF == PF_DEPTH32F) depth_.assign(w * h, 0.0f);
    }
    size_t width() const { return width_; }
    size_t height() const { return height_; }
private:
    std::vector<color_type> color_;
    std::vector<float> depth_;
    size_t width_ = 0, height_ = 0;
};
template <pixel_format CF = PF_RGBA32F, pixel_format DF = PF_UNSPECIFIED>
using cpu_buffer_rt = simple_buffer_rt<CF, DF>;   // cpu_buffer_rt minus display_color_buffer() (GL: unavailable on this box)

// device memory; frame() is asynchronous on the library's stream, as cuda_sched's launch is
template <pixel_format CF = PF_RGBA32F, pixel_format DF = PF_UNSPECIFIED>
class gpu_buffer_rt
{
    static_assert(DF == PF_UNSPECIFIED || DF == PF_DEPTH32F, "the path writes DEPTH32F depth");
public:
    using color_type = typename pixel_traits<CF>::type;
    using depth_type = float;
    using ref_type = render_target_ref<CF, DF>;
    static const int memory = VSB_MEM_DEVICE;
    static const int color_format = pixel_traits<CF>::id;

    gpu_buffer_rt() = default;
    gpu_buffer_rt(gpu_buffer_rt const&) = delete;
    gpu_buffer_rt& operator=(gpu_buffer_rt const&) = delete;
    ~gpu_buffer_rt() { if (color_) vsb_device_free(color_); if (depth_) vsb_device_free(depth_); }

    color_type* color() { return color_; }                      // DEVICE pointer
    color_type const* color() const { return color_; }
    depth_type* depth() { return depth_; }                      // DEVICE pointer (null without a depth format)
    depth_type const* depth() const { return depth_; }
    ref_type ref() { return ref_type{ color_, depth_, int(width_), int(height_) }; }
    void clear_depth_buffer(float d = 1.0f)
    {
        if (!depth_) return;
        std::vector<float> tmp(width_ * height_, d);
        detail::check(vsb_device_copy(depth_, tmp.data(), tmp.size() * sizeof(float), VSB_MEM_DEVICE, VSB_MEM_HOST), "vsb_device_copy");
    }
    void download_depth(float* dst) const
    {
        detail::check(vsb_device_copy(dst, depth_, width_ * height_ * sizeof(float), VSB_MEM_HOST, VSB_MEM_DEVICE), "vsb_device_copy");
    }
    void clear_color_buffer(vec4 const& c = vec4(0.0f))
    {
        std::vector<color_type> tmp(width_ * height_, detail::to_pixel<color_type>(c));
        detail::check(vsb_device_copy(color_, tmp.data(), tmp.size() * sizeof(color_type), VSB_MEM_DEVICE, VSB_MEM_HOST), "vsb_device_copy");
    }
    void begin_frame() {}
    void end_frame() {}
    void resize(size_t w, size_t h)
    {
        detail::ensure_init();
        if (color_) { vsb_device_free(color_); color_ = nullptr; }
        width_ = w; height_ = h;
        void* p = nullptr;
        detail::check(vsb_device_alloc(w * h * sizeof(color_type), &p), "vsb_device_alloc");
        color_ = static_cast<color_type*>(p);
        clear_color_buffer();
        if (depth_) { vsb_device_free(depth_); depth_ = nullptr; }
        if (DF == PF_DEPTH32F)
        {
            detail::check(vsb_device_alloc(w * h * sizeof(float), &p), "vsb_device_alloc");
            depth_ = static_cast<float*>(p);
            clear_depth_buffer(0.0f);
        }
    }
    size_t width() const { return width_; }
    size_t height() const { return height_; }
    // read back (the reference's users go through GL or thrust copies)
    void download(color_type* dst) const
    {
        detail::check(vsb_device_copy(dst, color_, width_ * height_ * sizeof(color_type), VSB_MEM_HOST, VSB_MEM_DEVICE), "vsb_device_copy");
    }
private:
    color_type* color_ = nullptr;
    float* depth_ = nullptr;
    size_t width_ = 0, height_ = 0;
};

//-------------------------------------------------------------------------------------------------
// Scheduler: cuda_sched<R> (detail/cuda_sched.h:24-40).  frame() calls rt.begin_frame() first and
// rt.end_frame() last (detail/cuda_sched.inl:309,319); kernel and sched_params are passed by
// value; not re-entrant.
//

template <typename R = ray>
class b200_sched
{
public:
    b200_sched() = default;
    explicit b200_sched(vec2ui) {}                    // block dims of cuda_sched: grids are sized by the library
    b200_sched(unsigned, unsigned) {}

    // image-space partition for multi-GPU: this process renders tiles k with k % nranks == rank
    void set_tile_partition(int rank, int nranks) { tile_rank_ = rank; tile_nranks_ = nranks; }
    void set_seed(unsigned seed) { seed_ = seed; }
    void set_reference_order(bool on) { traverse_ = on ? VSB_TRAVERSE_EXACT : VSB_TRAVERSE_WIDE; }
    // the surface arrays of the kernel params were modified in place: the next frame() uploads them again
    void surface_changed() { surface_print_.clear(); }

    template <typename K, typename SP>
    void frame(K kernel, SP sparams, unsigned frame_num = 0)
    {
        detail::ensure_init();
        sparams.rt.begin_frame();

        auto const& kp = kernel.params;
        vsb_frame_desc d; std::memset(&d, 0, sizeof(d));
        d.p.kernel = K::id;
        d.p.pixel_sampler = SP::pixel_sampler_type::id;
        d.p.width = int(sparams.rt.width()); d.p.height = int(sparams.rt.height());
        set_camera(d, sparams);
        // jittered_blend: alpha = 1 / frame_num (detail/sched_common.h:516-528); the reference's
        // callers count frames from 1 there
        d.p.frame_num = frame_num; d.p.seed = seed_;
        d.p.num_bounces = kp.num_bounces; d.p.epsilon = kp.epsilon;
        std::memcpy(d.p.bg, kp.bg_color.data(), 16); std::memcpy(d.p.ambient, kp.ambient_color.data(), 16);
        set_ao(d, kernel);
        std::vector<vsb_material> mats(kp.num_materials);
        for (size_t i = 0; i < kp.num_materials; ++i) mats[i] = kp.materials[i].flatten();
        std::vector<vsb_point_light> const& lights = kp.lights;
        d.p.n_materials = int(mats.size()); d.p.n_lights = int(lights.size());
        if (d.camera_kind == VSB_CAM_BASIS) detail::check(vsb_camera_basis(&d), "vsb_camera_basis");
        d.scissor[0] = sparams.scissor_box.x; d.scissor[1] = sparams.scissor_box.y;
        d.scissor[2] = sparams.scissor_box.w; d.scissor[3] = sparams.scissor_box.h;
        d.normals = reinterpret_cast<float const*>(kp.normals);
        d.materials = mats.data(); d.lights = lights.data();
        d.params_mem = VSB_MEM_HOST;
        d.color = sparams.rt.color();
        d.depth = d.camera_kind == VSB_CAM_MATRIX ? sparams.rt.depth() : nullptr;   // the reference's depth path needs the matrices
        d.mem = std::remove_reference<decltype(sparams.rt)>::type::memory;
        d.color_format = std::remove_reference<decltype(sparams.rt)>::type::color_format;
        d.traverse = traverse_;
        d.tile_rank = tile_rank_; d.tile_nranks = tile_nranks_;

        std::vector<vsb_scene const*> scenes;
        for (auto it = kp.prims_begin; it != kp.prims_end; ++it) scenes.push_back(it->scene);
        if (kp.surface.present && (kp.surface.normals || kp.surface.normal_binding == VSB_NORMALS_FROM_PRIMITIVE) && !scenes.empty())
        {
            // The arrays are deep-copied to the device (O(scene) upload + device syncs), so not on every frame: only when
            // the description handed in differs from the one this scheduler last attached to this scene -- array
            // addresses, bindings, texture descriptors.  A caller who rewrites the arrays IN PLACE between frames
            // calls surface_changed() (or b200_index_bvh::set_surface() once, which bypasses this path).
            std::vector<vsb_texture> texs;
            vsb_surface_desc sd = kp.surface.flatten(texs);
            std::vector<unsigned char> print(sizeof(void*) + sizeof(sd) + texs.size() * sizeof(vsb_texture));
            void const* scene0 = scenes[0];
            sd.textures = nullptr;                            // `texs` is a temporary: its address is not part of the identity
            std::memcpy(print.data(), &scene0, sizeof(void*));
            std::memcpy(print.data() + sizeof(void*), &sd, sizeof(sd));
            if (!texs.empty()) std::memcpy(print.data() + sizeof(void*) + sizeof(sd), texs.data(), texs.size() * sizeof(vsb_texture));
            sd.textures = texs.data();
            if (print != surface_print_)
            {
                detail::check(vsb_scene_set_surface(const_cast<vsb_scene*>(scenes[0]), &sd, VSB_MEM_HOST), "vsb_scene_set_surface");
                surface_print_.swap(print);
            }
        }
        detail::check(vsb_frame(scenes.data(), int(scenes.size()), &d, nullptr), "vsb_frame");

        sparams.rt.end_frame();
    }

private:
    template <typename RT, typename PS> static void set_camera(vsb_frame_desc& d, sched_params<RT, PS> const& sp)
    {
        for (int i = 0; i < 3; ++i) { d.p.eye[i] = sp.cam.eye()[i]; d.p.center[i] = sp.cam.center()[i]; d.p.up[i] = sp.cam.up()[i]; }
        d.p.fovy = sp.cam.fovy(); d.p.aspect = sp.cam.aspect();
        d.camera_kind = VSB_CAM_BASIS;
    }
    template <typename RT, typename PS> static void set_camera(vsb_frame_desc& d, sched_params_matrix<RT, PS> const& sp)
    {
        d.camera_kind = VSB_CAM_MATRIX;
        std::memcpy(d.view_matrix, sp.view_matrix.data(), 64); std::memcpy(d.proj_matrix, sp.proj_matrix.data(), 64);
    }
    template <typename K> static void set_ao(vsb_frame_desc&, K const&) {}
    template <typename KP> static void set_ao(vsb_frame_desc& d, ao::kernel<KP> const& k) { d.p.ao_samples = k.num_samples; d.p.ao_radius = k.radius; }

    int tile_rank_ = 0, tile_nranks_ = 0, traverse_ = VSB_TRAVERSE_WIDE;
    unsigned seed_ = 1;
    std::vector<unsigned char> surface_print_;     // identity of the surface description last attached by frame()
};

} // namespace visionaray_b200

#endif // VISIONARAY_B200_SCHEDULER_H
