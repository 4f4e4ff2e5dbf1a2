// TEST / BASELINE INFRASTRUCTURE -- not part of the product path.
//
// Thin extern "C" shim around the UNMODIFIED reference headers (tu500/visionaray,
// /root/reference/include).  It is compiled in the build container only (the GPU box has no
// /root/reference) by oracle/Makefile into oracle/_ref/libvsnray_ref*.so, which is
// git-ignored but travels to the GPU box.  Nothing from the reference is copied into this
// repository: this file only *calls* the reference's public API:
//   build<index_bvh<P>>()                 bvh.h:455, detail/bvh/build.inl:165-178
//   intersect<Traversal>(ray,bvh,...)     detail/bvh/intersect.inl:35-134
//   closest_hit()/any_hit()               detail/traverse_linear.inl:186-318
//   simple/whitted/pathtracing::kernel    detail/{simple,whitted,pathtracing}.inl
//   tiled_sched<R>::frame()               detail/tiled_sched.inl:365-391
//   make_primary_ray_impl                 detail/sched_common.h:130-150
//
// Two builds of this file exist (see Makefile):
//   libvsnray_ref.so       -O2 -ffp-contract=off  : the PARITY reference (no FMA contraction,
//                                                    so results do not depend on the compiler's
//                                                    contraction choices)
//   libvsnray_ref_fast.so  -O3 -mavx2 -mfma       : the TIMED CPU baseline (tiled_sched<ray8>)

#include <cstdint>
#include <cstring>
#include <cstdio>
#include <atomic>
#include <thread>
#include <vector>
#include <chrono>

#include <visionaray/math/math.h>
#include <visionaray/bvh.h>
#include <visionaray/traverse.h>
#include <visionaray/kernels.h>
#include <visionaray/material.h>
#include <visionaray/point_light.h>
#include <visionaray/spot_light.h>
#include <visionaray/scheduler.h>
#include <visionaray/simple_buffer_rt.h>
#include <visionaray/camera.h>
#include <visionaray/detail/tiled_sched.h>
#include <visionaray/sampling.h>
#include <visionaray/get_normal.h>
#include <visionaray/generic_material.h>
#include <visionaray/generic_primitive.h>
#include <visionaray/texture/texture.h>

using namespace visionaray;

namespace
{

using tri_t     = basic_triangle<3, float>;
using sph_t     = basic_sphere<float>;
using ray_t     = basic_ray<float>;
using tri_ref_t = index_bvh<tri_t>::bvh_ref;
using sph_ref_t = index_bvh<sph_t>::bvh_ref;
using gen_t     = generic_primitive<tri_t, sph_t>;      // 80 bytes: the alternative in place, its 1-based index at +64

// Layout shared with include/vsnray_b200.h (vsb_hit).
struct hit32
{
    int32_t  hit;
    uint32_t prim_id;
    uint32_t geom_id;
    float    t;
    float    u;
    float    v;
    uint32_t prim_list_index;
    uint32_t pad;
};

// Layout shared with include/vsnray_b200.h (vsb_material / vsb_point_light).
struct mat_pod
{
    float ca[3]; float ka;
    float cd[3]; float kd;
    float cs[3]; float ks;
    float spec_exp; int32_t kind; float pad[2];   // kind: 0 plastic, 1 matte
};

struct light_pod
{
    float pos[3]; float kl;
    float cl[3];  float att_const;
    float att_lin; float att_quad; int32_t kind; float spot_exponent;   // kind: 0 point_light, 1 spot_light
    float spot_dir[3]; float spot_cos_cutoff;
};

// Deterministic counter-based sampler (PCG-RXS-M-XS-32); the same generator is restated in
// oracle/vsnray_oracle.c and in the CUDA kernels.  The reference's schedulers hard-wire a
// time-seeded random_sampler (detail/tiled_sched.inl:191), which cannot be compared.
struct det_sampler
{
    using value_type = float;
    uint32_t s;
    explicit det_sampler(uint32_t pixel, uint32_t frame, uint32_t seed)
    {
        uint32_t h = seed * 0x9E3779B9u + 0x7F4A7C15u;
        h ^= pixel + 0x85EBCA6Bu + (h << 6) + (h >> 2);
        h ^= frame + 0xC2B2AE35u + (h << 6) + (h >> 2);
        h ^= h >> 16; h *= 0x7FEB352Du; h ^= h >> 15; h *= 0x846CA68Bu; h ^= h >> 16;
        s = h;
    }
    float next()
    {
        s = s * 747796405u + 2891336453u;
        uint32_t w = ((s >> ((s >> 28u) + 4u)) ^ s) * 277803737u;
        w = (w >> 22u) ^ w;
        return float(w >> 8) * (1.0f / 16777216.0f);
    }
};

template <typename HR>
inline void store_hit(hit32& o, HR const& hr)
{
    o.hit = hr.hit ? 1 : 0;
    o.prim_id = hr.hit ? (uint32_t)hr.prim_id : 0u;
    o.geom_id = hr.hit ? (uint32_t)hr.geom_id : 0u;
    o.t = hr.t;
    o.u = hr.hit ? hr.u : 0.0f;
    o.v = hr.hit ? hr.v : 0.0f;
    o.prim_list_index = hr.hit ? (uint32_t)hr.primitive_list_index : 0u;
    o.pad = 0;
}

template <typename F>
void parallel_for(size_t n, int nthreads, F f)
{
    if (nthreads <= 1)
    {
        f(0, n);
        return;
    }
    std::atomic<size_t> next(0);
    size_t const chunk = 4096;
    std::vector<std::thread> ts;
    for (int i = 0; i < nthreads; ++i)
    {
        ts.emplace_back([&]()
        {
            for (;;)
            {
                size_t b = next.fetch_add(chunk);
                if (b >= n) break;
                size_t e = b + chunk < n ? b + chunk : n;
                f(b, e);
            }
        });
    }
    for (auto& t : ts) t.join();
}

template <typename P>
typename index_bvh<P>::bvh_ref make_ref(
        void const* prims, size_t n_prims, void const* nodes, size_t n_nodes,
        uint32_t const* indices, size_t n_idx)
{
    using ref_t = typename index_bvh<P>::bvh_ref;
    auto p = reinterpret_cast<P const*>(prims);
    auto n = reinterpret_cast<bvh_node const*>(nodes);
    return ref_t(p, p + n_prims, n, n + n_nodes, indices, indices + n_idx);
}

template <typename P>
int build_impl(void const* prims, size_t n, int spatial_splits,
               void* nodes_out, size_t nodes_cap, size_t* n_nodes,
               uint32_t* idx_out, size_t idx_cap, size_t* n_idx)
{
    auto tree = build<index_bvh<P>>(reinterpret_cast<P const*>(prims), n, spatial_splits != 0);
    *n_nodes = tree.nodes().size();
    *n_idx   = tree.indices().size();
    if (*n_nodes > nodes_cap || *n_idx > idx_cap) return 1;
    std::memcpy(nodes_out, tree.nodes().data(), *n_nodes * sizeof(bvh_node));
    std::memcpy(idx_out, tree.indices().data(), *n_idx * sizeof(uint32_t));
    return 0;
}

template <typename P>
void query_impl(int any, void const* prims, size_t n_prims, void const* nodes, size_t n_nodes,
                uint32_t const* indices, size_t n_idx, void const* rays, size_t n_rays,
                float max_t, float const* max_t_per_ray, hit32* out, int nthreads)
{
    auto ref = make_ref<P>(prims, n_prims, nodes, n_nodes, indices, n_idx);
    auto r = reinterpret_cast<ray_t const*>(rays);
    parallel_for(n_rays, nthreads, [&](size_t b, size_t e)
    {
        for (size_t i = b; i < e; ++i)
        {
            float mt = max_t_per_ray ? max_t_per_ray[i] : max_t;
            default_intersector isect;
            if (any)
            {
                auto hr = any_hit(r[i], &ref, &ref + 1, mt, isect);
                store_hit(out[i], hr);
            }
            else
            {
                auto hr = intersect<detail::ClosestHit>(r[i], ref, isect, mt);
                store_hit(out[i], hr);
            }
        }
    });
}

// multi_hit<N>(r, begin, end) from the unmodified headers (detail/traverse_linear.inl:325-380),
// over a range of one or two bvh refs
template <typename P, size_t N>
void multi_impl(void const* prims, size_t n_prims, void const* nodes, size_t n_nodes, uint32_t const* indices, size_t n_idx,
                void const* prims2, size_t n_prims2, void const* nodes2, size_t n_nodes2, uint32_t const* indices2, size_t n_idx2,
                void const* rays, size_t n_rays, hit32* out, int nthreads)
{
    using ref_t = typename index_bvh<P>::bvh_ref;
    std::vector<ref_t> refs;
    refs.push_back(make_ref<P>(prims, n_prims, nodes, n_nodes, indices, n_idx));
    if (prims2) refs.push_back(make_ref<P>(prims2, n_prims2, nodes2, n_nodes2, indices2, n_idx2));
    auto r = reinterpret_cast<ray_t const*>(rays);
    parallel_for(n_rays, nthreads, [&](size_t b, size_t e)
    {
        for (size_t i = b; i < e; ++i)
        {
            auto hr = multi_hit<N>(r[i], refs.data(), refs.data() + refs.size());
            for (size_t k = 0; k < N; ++k) store_hit(out[i * N + k], hr[k]);
        }
    });
}

template <typename P>
int multi_dispatch(int n, void const* prims, size_t n_prims, void const* nodes, size_t n_nodes, uint32_t const* indices, size_t n_idx,
                   void const* prims2, size_t n_prims2, void const* nodes2, size_t n_nodes2, uint32_t const* indices2, size_t n_idx2,
                   void const* rays, size_t n_rays, hit32* out, int nthreads)
{
#define VSB_MULTI_CASE(N) case N: multi_impl<P, N>(prims, n_prims, nodes, n_nodes, indices, n_idx, prims2, n_prims2, nodes2, n_nodes2, indices2, n_idx2, rays, n_rays, out, nthreads); return 0;
    switch (n)
    {
    VSB_MULTI_CASE(1) VSB_MULTI_CASE(2) VSB_MULTI_CASE(4) VSB_MULTI_CASE(5) VSB_MULTI_CASE(8) VSB_MULTI_CASE(16)
    default: return 1;
    }
#undef VSB_MULTI_CASE
}

} // namespace

extern "C"
{

// n in {1, 2, 4, 5, 8, 16}; the second bvh (prims2 ...) may be NULL.  hits_out: n_rays * n records.
int ref_multi_hit(int prim_kind, int n,
                  void const* prims, size_t n_prims, void const* nodes, size_t n_nodes, uint32_t const* indices, size_t n_idx,
                  void const* prims2, size_t n_prims2, void const* nodes2, size_t n_nodes2, uint32_t const* indices2, size_t n_idx2,
                  void const* rays, size_t n_rays, void* hits_out, int nthreads)
{
    if (prim_kind == 2)
        return multi_dispatch<gen_t>(n, prims, n_prims, nodes, n_nodes, indices, n_idx, prims2, n_prims2, nodes2, n_nodes2, indices2, n_idx2, rays, n_rays, (hit32*)hits_out, nthreads);
    if (prim_kind == 0)
        return multi_dispatch<tri_t>(n, prims, n_prims, nodes, n_nodes, indices, n_idx, prims2, n_prims2, nodes2, n_nodes2, indices2, n_idx2, rays, n_rays, (hit32*)hits_out, nthreads);
    return multi_dispatch<sph_t>(n, prims, n_prims, nodes, n_nodes, indices, n_idx, prims2, n_prims2, nodes2, n_nodes2, indices2, n_idx2, rays, n_rays, (hit32*)hits_out, nthreads);
}

// sizes: {vec3, triangle, sphere, ray, bvh_node, aabb, off(tri.v1), off(tri.e1), off(tri.e2),
//         off(tri.prim_id), off(tri.geom_id), off(sphere.center), off(sphere.radius)}
void ref_layout(int* out)
{
    out[0] = (int)sizeof(vec3);
    out[1] = (int)sizeof(tri_t);
    out[2] = (int)sizeof(sph_t);
    out[3] = (int)sizeof(ray_t);
    out[4] = (int)sizeof(bvh_node);
    out[5] = (int)sizeof(aabb);
    tri_t t; sph_t s;
    out[6]  = (int)((char*)&t.v1 - (char*)&t);
    out[7]  = (int)((char*)&t.e1 - (char*)&t);
    out[8]  = (int)((char*)&t.e2 - (char*)&t);
    out[9]  = (int)((char*)&t.prim_id - (char*)&t);
    out[10] = (int)((char*)&t.geom_id - (char*)&t);
    out[11] = (int)((char*)&s.center - (char*)&s);
    out[12] = (int)((char*)&s.radius - (char*)&s);
}

int ref_build_triangles(void const* prims, size_t n, int spatial_splits,
                        void* nodes_out, size_t nodes_cap, size_t* n_nodes,
                        uint32_t* idx_out, size_t idx_cap, size_t* n_idx)
{
    return build_impl<tri_t>(prims, n, spatial_splits, nodes_out, nodes_cap, n_nodes, idx_out, idx_cap, n_idx);
}

int ref_build_spheres(void const* prims, size_t n, int spatial_splits,
                      void* nodes_out, size_t nodes_cap, size_t* n_nodes,
                      uint32_t* idx_out, size_t idx_cap, size_t* n_idx)
{
    return build_impl<sph_t>(prims, n, spatial_splits, nodes_out, nodes_cap, n_nodes, idx_out, idx_cap, n_idx);
}

int ref_build_generic(void const* prims, size_t n, int spatial_splits,
                      void* nodes_out, size_t nodes_cap, size_t* n_nodes,
                      uint32_t* idx_out, size_t idx_cap, size_t* n_idx)
{
    return build_impl<gen_t>(prims, n, spatial_splits, nodes_out, nodes_cap, n_nodes, idx_out, idx_cap, n_idx);
}

// generic_primitive<triangle, sphere> records from a tag array (0 triangle, 1 sphere) and the two payloads
// (64-byte triangle / 48-byte sphere per entry, the unused one ignored): the reference's own constructors
// decide the memory layout.  out: n * sizeof(gen_t) bytes.  Returns sizeof(gen_t).
int ref_make_generic(int const* tags, void const* tris, void const* sphs, size_t n, void* out)
{
    auto t = reinterpret_cast<tri_t const*>(tris);
    auto s = reinterpret_cast<sph_t const*>(sphs);
    if (out)
    {
        std::memset(out, 0, n * sizeof(gen_t));
        auto o = reinterpret_cast<gen_t*>(out);
        for (size_t i = 0; i < n; ++i) { if (tags[i] == 0) o[i] = gen_t(t[i]); else o[i] = gen_t(s[i]); }
    }
    return int(sizeof(gen_t));
}

// prim_kind: 0 triangles, 1 spheres, 2 generic_primitive<triangle, sphere>.  any: 0 closest-hit, 1 any-hit.
void ref_query(int prim_kind, int any,
               void const* prims, size_t n_prims, void const* nodes, size_t n_nodes,
               uint32_t const* indices, size_t n_idx, void const* rays, size_t n_rays,
               float max_t, float const* max_t_per_ray, void* hits_out, int nthreads)
{
    if (prim_kind == 2)
        query_impl<gen_t>(any, prims, n_prims, nodes, n_nodes, indices, n_idx, rays, n_rays, max_t, max_t_per_ray, (hit32*)hits_out, nthreads);
    else if (prim_kind == 0)
        query_impl<tri_t>(any, prims, n_prims, nodes, n_nodes, indices, n_idx, rays, n_rays, max_t, max_t_per_ray, (hit32*)hits_out, nthreads);
    else
        query_impl<sph_t>(any, prims, n_prims, nodes, n_nodes, indices, n_idx, rays, n_rays, max_t, max_t_per_ray, (hit32*)hits_out, nthreads);
}

} // extern "C"

#include "ref_harness_frame.inl"
