// TEST / BASELINE INFRASTRUCTURE -- frame() half of oracle/ref_harness.cpp.
// Calls the reference's kernel functors and schedulers; copies nothing from them.

namespace
{

// Mirrors vsb_frame_params in include/vsnray_b200.h (field for field).
struct frame_params
{
    int32_t  kernel;          // 0 simple, 1 whitted, 2 ao, 3 pathtracing
    int32_t  pixel_sampler;   // 0 uniform, 1 jittered, 2 jittered_blend
    int32_t  width, height;
    float    eye[4], center[4], up[4];
    float    fovy, aspect;    // radians
    uint32_t frame_num, seed;
    uint32_t num_bounces;
    float    epsilon;
    float    bg[4];
    float    ambient[4];
    int32_t  ao_samples;
    float    ao_radius;
    int32_t  n_lights, n_materials;
};

struct scene_host
{
    tri_ref_t ref;
    aligned_vector<vec3> normals;
    aligned_vector<plastic<float>> plastics;
    aligned_vector<matte<float>> mattes;
    aligned_vector<point_light<float>> lights;
};

inline spectrum<float> spec3(float const* c) { return spectrum<float>(from_rgb(c[0], c[1], c[2])); }

void fill_scene(scene_host& s, frame_params const& fp, float const* normals16, size_t n_normals,
                mat_pod const* mats, light_pod const* lights)
{
    s.normals.resize(n_normals);
    for (size_t i = 0; i < n_normals; ++i)
        s.normals[i] = vec3(normals16[i * 4], normals16[i * 4 + 1], normals16[i * 4 + 2]);
    for (int i = 0; i < fp.n_materials; ++i)
    {
        plastic<float> p;
        p.set_ca(spec3(mats[i].ca)); p.set_ka(mats[i].ka);
        p.set_cd(spec3(mats[i].cd)); p.set_kd(mats[i].kd);
        p.set_cs(spec3(mats[i].cs)); p.set_ks(mats[i].ks);
        p.set_specular_exp(mats[i].spec_exp);
        s.plastics.push_back(p);
        matte<float> m;
        m.set_ca(spec3(mats[i].ca)); m.set_ka(mats[i].ka);
        m.set_cd(spec3(mats[i].cd)); m.set_kd(mats[i].kd);
        s.mattes.push_back(m);
    }
    for (int i = 0; i < fp.n_lights; ++i)
    {
        point_light<float> l;
        l.set_cl(vec3(lights[i].cl[0], lights[i].cl[1], lights[i].cl[2]));
        l.set_kl(lights[i].kl);
        l.set_position(vec3(lights[i].pos[0], lights[i].pos[1], lights[i].pos[2]));
        l.set_constant_attenuation(lights[i].att_const);
        l.set_linear_attenuation(lights[i].att_lin);
        l.set_quadratic_attenuation(lights[i].att_quad);
        s.lights.push_back(l);
    }
}

camera make_cam(frame_params const& fp)
{
    camera cam;
    cam.perspective(fp.fovy, fp.aspect, 0.001f, 1000.0f);
    cam.look_at(vec3(fp.eye[0], fp.eye[1], fp.eye[2]),
                vec3(fp.center[0], fp.center[1], fp.center[2]),
                vec3(fp.up[0], fp.up[1], fp.up[2]));
    return cam;
}

// The AO kernel of src/examples/ao/main.cpp:183-246, instantiated through the reference's
// own closest_hit / any_hit / get_normal / make_orthonormal_basis / cosine_sample_hemisphere.
template <typename R>
struct ao_kernel
{
    using S = typename R::scalar_type;
    using C = vector<4, S>;
    using V = vector<3, S>;

    tri_ref_t const* begin;
    tri_ref_t const* end;
    vec3 const* normals;
    vec4 bg;
    int samples;
    float radius;

    template <typename Sampler>
    result_record<S> operator()(R ray, Sampler& samp) const
    {
        result_record<S> result;
        result.color = C(bg);
        auto hit_rec = closest_hit(ray, begin, end);
        result.hit = hit_rec.hit;
        if (any(hit_rec.hit))
        {
            hit_rec.isect_pos = ray.ori + ray.dir * hit_rec.t;
            result.isect_pos  = hit_rec.isect_pos;
            V clr(1.0);
            auto n = get_normal(normals, hit_rec, index_bvh<tri_t>{}, normals_per_face_binding{});
            V u, v, w = n;
            make_orthonormal_basis(u, v, w);
            S rad = radius;
            for (int i = 0; i < samples; ++i)
            {
                auto sp = cosine_sample_hemisphere(samp.next(), samp.next());
                auto dir = normalize(sp.x * u + sp.y * v + sp.z * w);
                R ao_ray;
                ao_ray.ori = hit_rec.isect_pos + dir * S(1E-3f);
                ao_ray.dir = dir;
                auto ao_rec = any_hit(ao_ray, begin, end, rad);
                clr = select(ao_rec.hit, clr - V(S(1.0f / samples)), clr);
            }
            result.color = select(hit_rec.hit, C(clr, S(1.0)), result.color);
        }
        return result;
    }
};

template <typename K>
struct no_sampler_adapter
{
    K k;
    template <typename R, typename Sampler>
    auto operator()(R ray, Sampler&) const -> decltype(k(ray)) { return k(ray); }
};

// Scalar per-pixel driver with the deterministic sampler: the PARITY frame oracle.
// matrix camera + depth target of the next scalar_frame() call (ref_frame_matrix sets them)
struct matrix_cam { bool on = false; mat4 view, inv_view, proj, inv_proj; float* depth = nullptr; };
static matrix_cam g_mcam;

template <typename K>
void scalar_frame(K kernel, frame_params const& fp, float* rgba, int nthreads, float* basis_out)
{
    matrix_cam const mc = g_mcam;
    auto cam = make_cam(fp);
    // host-side basis, as the reference schedulers compute it (detail/tiled_sched.inl:296-304)
    auto f = normalize(cam.eye() - cam.center());
    auto s = normalize(cross(cam.up(), f));
    auto u = cross(f, s);
    vec3 eye   = cam.eye();
    vec3 cam_u = s * tan(cam.fovy() / 2.0f) * cam.aspect();
    vec3 cam_v = u * tan(cam.fovy() / 2.0f);
    vec3 cam_w = -f;
    if (basis_out)
    {
        float b[12] = { eye.x, eye.y, eye.z, cam_u.x, cam_u.y, cam_u.z, cam_v.x, cam_v.y, cam_v.z, cam_w.x, cam_w.y, cam_w.z };
        std::memcpy(basis_out, b, sizeof(b));
    }
    size_t w = fp.width, h = fp.height;
    parallel_for(w * h, nthreads, [&](size_t b, size_t e)
    {
        for (size_t i = b; i < e; ++i)
        {
            unsigned x = unsigned(i % w), y = unsigned(i / w);
            det_sampler samp((uint32_t)i, fp.frame_num, fp.seed);
            float px = float(x), py = float(y);
            float* dst = rgba + i * 4;
            if (fp.pixel_sampler >= 3)
            {
                // ssaa_type<N> with the reference's own ray generator: the N rays of
                // make_primary_rays(ssaa_type<N>) (detail/sched_common.h:222-295), then clear +
                // N x blend(src / N + dst) (detail/sched_common.h:640-698)
                vec4 acc(0.0f);
                float zacc = 0.0f;
                auto run = [&](ray_t const& r, float alpha)
                {
                    auto res = kernel(r, samp);
                    acc = res.color * alpha + acc * 1.0f;
                    if (mc.on)
                    {
                        float z = res.hit ? detail::depth_transform(res.isect_pos, mc.view, mc.inv_view, mc.proj, mc.inv_proj) : 1.0f;
                        zacc = z * alpha + zacc * 1.0f;
                    }
                };
                random_sampler<float> unused;
                if (mc.on)
                {
                    // the same generators with the matrix argument pack (detail/sched_common.h:152-178)
                    if (fp.pixel_sampler == 3)
                    {
                        auto rays = detail::make_primary_rays(ray_t{}, pixel_sampler::ssaa_type<2>{}, unused, x, y, w, h, mc.view, mc.inv_view, mc.proj, mc.inv_proj);
                        for (auto const& r : rays) run(r, 1.0f / 2.0f);
                    }
                    else if (fp.pixel_sampler == 4)
                    {
                        auto rays = detail::make_primary_rays(ray_t{}, pixel_sampler::ssaa_type<4>{}, unused, x, y, w, h, mc.view, mc.inv_view, mc.proj, mc.inv_proj);
                        for (auto const& r : rays) run(r, 1.0f / 4.0f);
                    }
                    else
                    {
                        auto rays = detail::make_primary_rays(ray_t{}, pixel_sampler::ssaa_type<8>{}, unused, x, y, w, h, mc.view, mc.inv_view, mc.proj, mc.inv_proj);
                        for (auto const& r : rays) run(r, 1.0f / 8.0f);
                    }
                    if (mc.depth) mc.depth[i] = zacc;
                }
                else if (fp.pixel_sampler == 3)
                {
                    auto rays = detail::make_primary_rays(ray_t{}, pixel_sampler::ssaa_type<2>{}, unused, x, y, w, h, eye, cam_u, cam_v, cam_w);
                    for (auto const& r : rays) run(r, 1.0f / 2.0f);
                }
                else if (fp.pixel_sampler == 4)
                {
                    auto rays = detail::make_primary_rays(ray_t{}, pixel_sampler::ssaa_type<4>{}, unused, x, y, w, h, eye, cam_u, cam_v, cam_w);
                    for (auto const& r : rays) run(r, 1.0f / 4.0f);
                }
                else
                {
                    auto rays = detail::make_primary_rays(ray_t{}, pixel_sampler::ssaa_type<8>{}, unused, x, y, w, h, eye, cam_u, cam_v, cam_w);
                    for (auto const& r : rays) run(r, 1.0f / 8.0f);
                }
                dst[0] = acc.x; dst[1] = acc.y; dst[2] = acc.z; dst[3] = acc.w;
                continue;
            }
            if (fp.pixel_sampler != 0)
            {
                // jittered: detail/sched_common.h:197-217
                float jx = samp.next() - 0.5f;
                float jy = samp.next() - 0.5f;
                px += jx; py += jy;
            }
            ray_t r = mc.on ? detail::make_primary_ray_impl<ray_t>(px, py, w, h, mc.view, mc.inv_view, mc.proj, mc.inv_proj)
                            : detail::make_primary_ray_impl<ray_t>(px, py, w, h, eye, cam_u, cam_v, cam_w);
            auto res = kernel(r, samp);
            vec4 c = res.color;
            if (mc.on && mc.depth)
            {
                // result.depth (detail/sched_common.h:425, :559), stored or blended with the colour
                float z = res.hit ? detail::depth_transform(res.isect_pos, mc.view, mc.inv_view, mc.proj, mc.inv_proj) : 1.0f;
                if (fp.pixel_sampler == 2)
                {
                    float alpha = 1.0f / float(fp.frame_num);
                    mc.depth[i] = z * alpha + mc.depth[i] * (1.0f - alpha);
                }
                else mc.depth[i] = z;
            }
            if (fp.pixel_sampler == 2)
            {
                // jittered_blend: detail/sched_common.h:516-528
                float alpha = 1.0f / float(fp.frame_num);
                dst[0] = c.x * alpha + dst[0] * (1.0f - alpha);
                dst[1] = c.y * alpha + dst[1] * (1.0f - alpha);
                dst[2] = c.z * alpha + dst[2] * (1.0f - alpha);
                dst[3] = c.w * alpha + dst[3] * (1.0f - alpha);
            }
            else
            {
                dst[0] = c.x; dst[1] = c.y; dst[2] = c.z; dst[3] = c.w;
            }
        }
    });
}

// tiled_sched<R>::frame() driver: the TIMED CPU baseline (and, for RNG-free kernels with the
// uniform pixel sampler, a second frame oracle).  Returns seconds for `reps` frames (best).
template <typename R, typename K, typename PS>
double tiled_frame(K kernel, PS ps, frame_params const& fp, float* rgba, int nthreads, int reps, int warmup)
{
    static tiled_sched<R>* sched = nullptr;
    static int sched_threads = 0;
    if (!sched || sched_threads != nthreads)
    {
        delete sched;
        sched = new tiled_sched<R>((unsigned)nthreads);
        sched_threads = nthreads;
        // workers must reach their condvar wait before the first notify
        std::this_thread::sleep_for(std::chrono::milliseconds(200));
    }
    simple_buffer_rt<PF_RGBA32F, PF_UNSPECIFIED> rt;
    rt.resize(fp.width, fp.height);
    auto cam = make_cam(fp);
    cam.set_viewport(0, 0, fp.width, fp.height);
    auto sparams = make_sched_params(ps, cam, rt);
    double best = 1e30;
    for (int i = 0; i < warmup + reps; ++i)
    {
        auto t0 = std::chrono::steady_clock::now();
        sched->frame(kernel, sparams, fp.frame_num);
        auto t1 = std::chrono::steady_clock::now();
        double dt = std::chrono::duration<double>(t1 - t0).count();
        if (i >= warmup && dt < best) best = dt;
    }
    if (rgba) std::memcpy(rgba, rt.color(), sizeof(float) * 4 * size_t(fp.width) * fp.height);
    return best;
}

// uniform or ssaa_type<N> by fp.pixel_sampler (RNG-free kernels: tiled_sched itself is the oracle)
template <typename R, typename K>
double tiled_frame_ps(K kernel, frame_params const& fp, float* rgba, int nthreads, int reps, int warmup)
{
    switch (fp.pixel_sampler)
    {
    case 3: return tiled_frame<R>(kernel, pixel_sampler::ssaa_type<2>{}, fp, rgba, nthreads, reps, warmup);
    case 4: return tiled_frame<R>(kernel, pixel_sampler::ssaa_type<4>{}, fp, rgba, nthreads, reps, warmup);
    case 5: return tiled_frame<R>(kernel, pixel_sampler::ssaa_type<8>{}, fp, rgba, nthreads, reps, warmup);
    default: return tiled_frame<R>(kernel, pixel_sampler::uniform_type{}, fp, rgba, nthreads, reps, warmup);
    }
}

template <typename R>
struct ao_tiled_adapter
{
    ao_kernel<R> k;
    template <typename Sampler>
    auto operator()(R ray, Sampler& s) const -> decltype(k(ray, s)) { return k(ray, s); }
};


// ---- kernel_params with a surface description: generic_material<emissive, matte,
// mirror, plastic>, normals per face / per vertex, optional colours (per face / per vertex),
// tex coords and texture_ref<vector<4, unorm<8>>, 2> -- all through the reference's own
// make_kernel_params / get_surface / tex2D.  Scalar parity driver only.

struct tex_pod { void const* data; int32_t width, height, filter, address[2]; };

using gen_mat_t = generic_material<emissive<float>, matte<float>, mirror<float>, plastic<float>>;
using tex_ref_t = texture_ref<vector<4, unorm<8>>, 2>;

gen_mat_t make_generic(mat_pod const& m)
{
    if (m.kind == 2)
    {
        emissive<float> e;
        e.set_ce(spec3(m.cd)); e.set_ls(m.kd);
        return gen_mat_t(e);
    }
    if (m.kind == 3)
    {
        mirror<float> r;
        r.set_cr(spec3(m.cs)); r.set_kr(m.ks);
        r.set_ior(spec3(m.ca)); r.set_absorption(spec3(m.cd));
        return gen_mat_t(r);
    }
    if (m.kind == 1)
    {
        matte<float> t;
        t.set_ca(spec3(m.ca)); t.set_ka(m.ka);
        t.set_cd(spec3(m.cd)); t.set_kd(m.kd);
        return gen_mat_t(t);
    }
    plastic<float> p;
    p.set_ca(spec3(m.ca)); p.set_ka(m.ka);
    p.set_cd(spec3(m.cd)); p.set_kd(m.kd);
    p.set_cs(spec3(m.cs)); p.set_ks(m.ks);
    p.set_specular_exp(m.spec_exp);
    return gen_mat_t(p);
}

template <typename KP>
void run_surface_kernels(KP kparams, frame_params const& fp, float* rgba, int nthreads)
{
    if (fp.kernel == 0)
    {
        simple::kernel<KP> k{ kparams };
        scalar_frame(no_sampler_adapter<simple::kernel<KP>>{ k }, fp, rgba, nthreads, nullptr);
    }
    else if (fp.kernel == 1)
    {
        whitted::kernel<KP> k{ kparams };
        scalar_frame(no_sampler_adapter<whitted::kernel<KP>>{ k }, fp, rgba, nthreads, nullptr);
    }
    else
    {
        pathtracing::kernel<KP> k{ kparams };
        scalar_frame(k, fp, rgba, nthreads, nullptr);
    }
}

// kernel_params WITHOUT normals (the default make_kernel_params overload, kernels.h:328-366): the normal comes from
// the primitive -- get_normal(hr, triangle | sphere | generic_primitive) -- for any primitive type
template <typename P>
int frame_prim_normals(frame_params const& fp, void const* prims, size_t n_prims, void const* nodes, size_t n_nodes,
                       uint32_t const* indices, size_t n_idx, void const* mats_, void const* lights_, float* rgba, int nthreads)
{
    auto ref = make_ref<P>(prims, n_prims, nodes, n_nodes, indices, n_idx);
    scene_host s{ tri_ref_t() };
    fill_scene(s, fp, nullptr, 0, (mat_pod const*)mats_, (light_pod const*)lights_);
    aligned_vector<gen_mat_t> mats;
    for (int i = 0; i < fp.n_materials; ++i) mats.push_back(make_generic(((mat_pod const*)mats_)[i]));
    vec4 bg(fp.bg[0], fp.bg[1], fp.bg[2], fp.bg[3]);
    vec4 amb(fp.ambient[0], fp.ambient[1], fp.ambient[2], fp.ambient[3]);
    auto lb = s.lights.data();
    auto le = s.lights.data() + s.lights.size();
    run_surface_kernels(make_kernel_params(&ref, &ref + 1, mats.data(), lb, le, fp.num_bounces, fp.epsilon, bg, amb),
                        fp, rgba, nthreads);
    return 0;
}

} // namespace

extern "C" int ref_frame_prim_normals(int prim_kind, void const* fp_, void const* prims, size_t n_prims, void const* nodes,
                                      size_t n_nodes, uint32_t const* indices, size_t n_idx, void const* mats_,
                                      void const* lights_, float* rgba, int nthreads)
{
    frame_params fp;
    std::memcpy(&fp, fp_, sizeof(fp));
    if (fp.kernel == 2) return -1;
    if (fp.kernel == 3 && prim_kind != 0) return -1;     // pathtracing.inl reads isect_pos before it is set: triangles only
    if (prim_kind == 0) return frame_prim_normals<tri_t>(fp, prims, n_prims, nodes, n_nodes, indices, n_idx, mats_, lights_, rgba, nthreads);
    if (prim_kind == 1) return frame_prim_normals<sph_t>(fp, prims, n_prims, nodes, n_nodes, indices, n_idx, mats_, lights_, rgba, nthreads);
    return -1;     // get_surface() does not instantiate for a generic_primitive without a normals array (get_surface.h:218-245)
}

namespace
{
} // namespace

// spot_light<float> list (spot_light.h): plastic materials, per-face normals, simple / whitted kernels.
// cutoff_angles: the angles the lights were set up with (the class keeps cos(cutoff) itself).
extern "C" int ref_frame_spot(void const* fp_, void const* prims, size_t n_prims, void const* nodes, size_t n_nodes,
                              uint32_t const* indices, size_t n_idx, float const* normals16, void const* mats_,
                              void const* lights_, float const* cutoff_angles, float* rgba, int nthreads)
{
    frame_params fp;
    std::memcpy(&fp, fp_, sizeof(fp));
    if (fp.kernel > 1) return -1;
    scene_host s{ make_ref<tri_t>(prims, n_prims, nodes, n_nodes, indices, n_idx) };
    fill_scene(s, fp, normals16, n_prims, (mat_pod const*)mats_, (light_pod const*)lights_);
    aligned_vector<spot_light<float>> spots;
    light_pod const* lp = (light_pod const*)lights_;
    for (int i = 0; i < fp.n_lights; ++i)
    {
        spot_light<float> l;
        l.set_cl(vec3(lp[i].cl[0], lp[i].cl[1], lp[i].cl[2]));
        l.set_kl(lp[i].kl);
        l.set_position(vec3(lp[i].pos[0], lp[i].pos[1], lp[i].pos[2]));
        l.set_spot_direction(vec3(lp[i].spot_dir[0], lp[i].spot_dir[1], lp[i].spot_dir[2]));
        l.set_spot_cutoff(cutoff_angles[i]);
        l.set_spot_exponent(lp[i].spot_exponent);
        l.set_constant_attenuation(lp[i].att_const);
        l.set_linear_attenuation(lp[i].att_lin);
        l.set_quadratic_attenuation(lp[i].att_quad);
        spots.push_back(l);
    }
    vec4 bg(fp.bg[0], fp.bg[1], fp.bg[2], fp.bg[3]);
    vec4 amb(fp.ambient[0], fp.ambient[1], fp.ambient[2], fp.ambient[3]);
    tri_ref_t const* pb = &s.ref;
    tri_ref_t const* pe = &s.ref + 1;
    auto kp = make_kernel_params(normals_per_face_binding{}, pb, pe, s.normals.data(), s.plastics.data(),
                                 spots.data(), spots.data() + spots.size(), fp.num_bounces, fp.epsilon, bg, amb);
    using KP = decltype(kp);
    if (fp.kernel == 0) scalar_frame(no_sampler_adapter<simple::kernel<KP>>{ { kp } }, fp, rgba, nthreads, nullptr);
    else                scalar_frame(no_sampler_adapter<whitted::kernel<KP>>{ { kp } }, fp, rgba, nthreads, nullptr);
    return 0;
}

// index_bvh<generic_primitive<triangle, sphere>> with a normals array (per face / per vertex): the visitor of
// detail/generic_primitive_get_normal.h picks the alternative's rule.  simple / whitted kernels.
extern "C" int ref_frame_generic(void const* fp_, void const* prims, size_t n_prims, void const* nodes, size_t n_nodes,
                                 uint32_t const* indices, size_t n_idx, int normal_binding, float const* normals16,
                                 void const* mats_, void const* lights_, float* rgba, int nthreads)
{
    frame_params fp;
    std::memcpy(&fp, fp_, sizeof(fp));
    if (fp.kernel > 1) return -1;
    auto ref = make_ref<gen_t>(prims, n_prims, nodes, n_nodes, indices, n_idx);
    scene_host s{ tri_ref_t() };
    fill_scene(s, fp, nullptr, 0, (mat_pod const*)mats_, (light_pod const*)lights_);
    aligned_vector<gen_mat_t> mats;
    for (int i = 0; i < fp.n_materials; ++i) mats.push_back(make_generic(((mat_pod const*)mats_)[i]));
    vec3 const* normals = reinterpret_cast<vec3 const*>(normals16);
    vec4 bg(fp.bg[0], fp.bg[1], fp.bg[2], fp.bg[3]);
    vec4 amb(fp.ambient[0], fp.ambient[1], fp.ambient[2], fp.ambient[3]);
    auto lb = s.lights.data();
    auto le = s.lights.data() + s.lights.size();
    if (normal_binding == 1)
        run_surface_kernels(make_kernel_params(normals_per_vertex_binding{}, &ref, &ref + 1, normals, mats.data(), lb, le,
                                               fp.num_bounces, fp.epsilon, bg, amb), fp, rgba, nthreads);
    else
        run_surface_kernels(make_kernel_params(normals_per_face_binding{}, &ref, &ref + 1, normals, mats.data(), lb, le,
                                               fp.num_bounces, fp.epsilon, bg, amb), fp, rgba, nthreads);
    return 0;
}

extern "C" void ref_tex2d(void const* tex_, float s, float t, float* out4)
{
    tex_pod const& tp = *(tex_pod const*)tex_;
    tex_ref_t tex(size_t(tp.width), size_t(tp.height));
    tex.reset((vector<4, unorm<8>> const*)tp.data);
    tex.set_filter_mode(tex_filter_mode(tp.filter));
    tex.set_address_mode(0, tex_address_mode(tp.address[0]));
    tex.set_address_mode(1, tex_address_mode(tp.address[1]));
    vec4 c = tex2D(tex, vec2(s, t));
    out4[0] = c.x; out4[1] = c.y; out4[2] = c.z; out4[3] = c.w;
}

extern "C" int ref_frame_surface(void const* fp_, void const* prims, size_t n_prims, void const* nodes, size_t n_nodes,
                                 uint32_t const* indices, size_t n_idx, int normal_binding, int color_binding,
                                 float const* normals16, float const* colors16, float const* tex_coords,
                                 void const* textures_, int n_textures, void const* mats_, void const* lights_,
                                 float* rgba, int nthreads)
{
    frame_params fp;
    std::memcpy(&fp, fp_, sizeof(fp));
    if (fp.kernel == 2) return -1;
    scene_host s{ make_ref<tri_t>(prims, n_prims, nodes, n_nodes, indices, n_idx) };
    fill_scene(s, fp, nullptr, 0, (mat_pod const*)mats_, (light_pod const*)lights_);
    aligned_vector<gen_mat_t> mats;
    for (int i = 0; i < fp.n_materials; ++i) mats.push_back(make_generic(((mat_pod const*)mats_)[i]));
    vec3 const* normals = reinterpret_cast<vec3 const*>(normals16);
    vec3 const* colors  = reinterpret_cast<vec3 const*>(colors16);
    // params with colours always carry tex coords + textures (kernels.h:271-319): without any, every
    // texture is empty (tex_color 1, get_surface.h:443-445) and the coordinates are never read
    aligned_vector<vec2> dummy_tc;
    vec2 const* tcs = reinterpret_cast<vec2 const*>(tex_coords);
    aligned_vector<tex_ref_t> texs;
    bool const with_tex = tex_coords != nullptr;
    if (!with_tex) { dummy_tc.resize(n_prims * 3, vec2(0.0f)); tcs = dummy_tc.data(); }
    for (int i = 0; i < fp.n_materials; ++i)
    {
        tex_ref_t t(0, 0);
        if (with_tex && i < n_textures)
        {
            tex_pod const& tp = ((tex_pod const*)textures_)[i];
            t = tex_ref_t(size_t(tp.width), size_t(tp.height));
            t.reset((vector<4, unorm<8>> const*)tp.data);
            t.set_filter_mode(tex_filter_mode(tp.filter));
            t.set_address_mode(0, tex_address_mode(tp.address[0]));
            t.set_address_mode(1, tex_address_mode(tp.address[1]));
        }
        texs.push_back(t);
    }
    vec4 bg(fp.bg[0], fp.bg[1], fp.bg[2], fp.bg[3]);
    vec4 amb(fp.ambient[0], fp.ambient[1], fp.ambient[2], fp.ambient[3]);
    auto lb = s.lights.data();
    auto le = s.lights.data() + s.lights.size();
    tri_ref_t const* pb = &s.ref;
    tri_ref_t const* pe = &s.ref + 1;

#define VSB_SURF(NB)                                                                                                         \
    do {                                                                                                                     \
        if (color_binding == 1)                                                                                              \
            run_surface_kernels(make_kernel_params(NB{}, colors_per_face_binding{}, pb, pe, normals, tcs, mats.data(), colors, \
                                                   texs.data(), lb, le, fp.num_bounces, fp.epsilon, bg, amb), fp, rgba, nthreads); \
        else if (color_binding == 2)                                                                                         \
            run_surface_kernels(make_kernel_params(NB{}, colors_per_vertex_binding{}, pb, pe, normals, tcs, mats.data(), colors, \
                                                   texs.data(), lb, le, fp.num_bounces, fp.epsilon, bg, amb), fp, rgba, nthreads); \
        else if (with_tex)                                                                                                   \
            run_surface_kernels(make_kernel_params(NB{}, pb, pe, normals, tcs, mats.data(), texs.data(), lb, le,             \
                                                   fp.num_bounces, fp.epsilon, bg, amb), fp, rgba, nthreads);                \
        else                                                                                                                 \
            run_surface_kernels(make_kernel_params(NB{}, pb, pe, normals, mats.data(), lb, le,                               \
                                                   fp.num_bounces, fp.epsilon, bg, amb), fp, rgba, nthreads);                \
    } while (0)
    if (normal_binding == 1) VSB_SURF(normals_per_vertex_binding);
    else                     VSB_SURF(normals_per_face_binding);
#undef VSB_SURF
    return 0;
}

namespace
{

// ---- render targets of other colour formats: the reference's own pixel_access::store / blend
// (detail/pixel_access.h) and colour conversion, driven per pixel like sample_pixel_impl does
// (detail/sched_common.h:361-528 store, :516-528 blend, :640-698 SSAA = clear + N x blend)

template <pixel_format CF, typename K>
void scalar_frame_format(K kernel, frame_params const& fp, void* color, int nthreads)
{
    using color_t = typename pixel_traits<CF>::type;
    color_t* buf = reinterpret_cast<color_t*>(color);
    auto cam = make_cam(fp);
    auto f = normalize(cam.eye() - cam.center());
    auto s = normalize(cross(cam.up(), f));
    auto u = cross(f, s);
    vec3 eye   = cam.eye();
    vec3 cam_u = s * tan(cam.fovy() / 2.0f) * cam.aspect();
    vec3 cam_v = u * tan(cam.fovy() / 2.0f);
    vec3 cam_w = -f;
    int w = fp.width, h = fp.height;
    pixel_format_constant<CF> dst_format;
    pixel_format_constant<PF_RGBA32F> src_format;
    parallel_for(size_t(w) * h, nthreads, [&](size_t b, size_t e)
    {
        for (size_t i = b; i < e; ++i)
        {
            int x = int(i % size_t(w)), y = int(i / size_t(w));
            det_sampler samp((uint32_t)i, fp.frame_num, fp.seed);
            float px = float(x), py = float(y);
            if (fp.pixel_sampler >= 3)
            {
                random_sampler<float> unused;
                detail::pixel_access::store(dst_format, src_format, x, y, w, h, vec4(0.0f), buf);
                auto run = [&](ray_t const& r, float alpha)
                {
                    auto res = kernel(r, samp);
                    detail::pixel_access::blend(dst_format, src_format, x, y, w, h, res.color, buf, alpha, 1.0f);
                };
                if (fp.pixel_sampler == 3)
                {
                    auto rays = detail::make_primary_rays(ray_t{}, pixel_sampler::ssaa_type<2>{}, unused, unsigned(x), unsigned(y), w, h, eye, cam_u, cam_v, cam_w);
                    for (auto const& r : rays) run(r, 1.0f / 2.0f);
                }
                else if (fp.pixel_sampler == 4)
                {
                    auto rays = detail::make_primary_rays(ray_t{}, pixel_sampler::ssaa_type<4>{}, unused, unsigned(x), unsigned(y), w, h, eye, cam_u, cam_v, cam_w);
                    for (auto const& r : rays) run(r, 1.0f / 4.0f);
                }
                else
                {
                    auto rays = detail::make_primary_rays(ray_t{}, pixel_sampler::ssaa_type<8>{}, unused, unsigned(x), unsigned(y), w, h, eye, cam_u, cam_v, cam_w);
                    for (auto const& r : rays) run(r, 1.0f / 8.0f);
                }
                continue;
            }
            if (fp.pixel_sampler != 0)
            {
                float jx = samp.next() - 0.5f;
                float jy = samp.next() - 0.5f;
                px += jx; py += jy;
            }
            ray_t r = detail::make_primary_ray_impl<ray_t>(px, py, w, h, eye, cam_u, cam_v, cam_w);
            auto res = kernel(r, samp);
            if (fp.pixel_sampler == 2)
            {
                float alpha = 1.0f / float(fp.frame_num);
                detail::pixel_access::blend(dst_format, src_format, x, y, w, h, res.color, buf, alpha, 1.0f - alpha);
            }
            else detail::pixel_access::store(dst_format, src_format, x, y, w, h, res.color, buf);
        }
    });
}

template <typename K>
int scalar_frame_any_format(K kernel, frame_params const& fp, void* color, int format, int nthreads)
{
    switch (format)
    {
    case 0: scalar_frame_format<PF_RGBA32F>(kernel, fp, color, nthreads); return 0;
    case 1: scalar_frame_format<PF_RGBA8>(kernel, fp, color, nthreads); return 0;
    case 2: scalar_frame_format<PF_RGB32F>(kernel, fp, color, nthreads); return 0;
    case 3: scalar_frame_format<PF_RGB8>(kernel, fp, color, nthreads); return 0;
    default: return -1;
    }
}

} // namespace

// scalar parity driver into a colour buffer of `format` (0 PF_RGBA32F, 1 PF_RGBA8, 2 PF_RGB32F, 3 PF_RGB8);
// plastic or matte material list, per-face normals, kernels simple / whitted / pathtracing
extern "C" int ref_frame_format(void const* fp_, void const* prims, size_t n_prims, void const* nodes, size_t n_nodes,
                                uint32_t const* indices, size_t n_idx, float const* normals16, void const* mats_,
                                void const* lights_, void* color, int format, int nthreads)
{
    frame_params fp;
    std::memcpy(&fp, fp_, sizeof(fp));
    if (fp.kernel == 2) return -1;
    scene_host s{ make_ref<tri_t>(prims, n_prims, nodes, n_nodes, indices, n_idx) };
    fill_scene(s, fp, normals16, n_prims, (mat_pod const*)mats_, (light_pod const*)lights_);
    bool use_matte = fp.n_materials > 0 && ((mat_pod const*)mats_)[0].kind == 1;
    vec4 bg(fp.bg[0], fp.bg[1], fp.bg[2], fp.bg[3]);
    vec4 amb(fp.ambient[0], fp.ambient[1], fp.ambient[2], fp.ambient[3]);
    auto lb = s.lights.data();
    auto le = s.lights.data() + s.lights.size();
    tri_ref_t const* pb = &s.ref;
    tri_ref_t const* pe = &s.ref + 1;
    auto run = [&](auto kparams) -> int
    {
        using KP = decltype(kparams);
        if (fp.kernel == 0) return scalar_frame_any_format(no_sampler_adapter<simple::kernel<KP>>{ { kparams } }, fp, color, format, nthreads);
        if (fp.kernel == 1) return scalar_frame_any_format(no_sampler_adapter<whitted::kernel<KP>>{ { kparams } }, fp, color, format, nthreads);
        return scalar_frame_any_format(pathtracing::kernel<KP>{ kparams }, fp, color, format, nthreads);
    };
    if (use_matte)
        return run(make_kernel_params(normals_per_face_binding{}, pb, pe, s.normals.data(), s.mattes.data(), lb, le,
                                      fp.num_bounces, fp.epsilon, bg, amb));
    return run(make_kernel_params(normals_per_face_binding{}, pb, pe, s.normals.data(), s.plastics.data(), lb, le,
                                  fp.num_bounces, fp.epsilon, bg, amb));
}

namespace
{
} // namespace

extern "C"
{

// mode 0: scalar parity driver (deterministic sampler) -> returns 0.0
// mode 1: tiled_sched<simd::ray8> (AVX2) timed; returns best seconds over reps
// mode 2: tiled_sched<ray> scalar timed
double ref_frame(void const* fp_, void const* prims, size_t n_prims, void const* nodes, size_t n_nodes,
                 uint32_t const* indices, size_t n_idx, float const* normals16,
                 void const* mats_, void const* lights_, float* rgba, int nthreads,
                 int mode, int reps, int warmup, float* basis_out)
{
    frame_params fp;
    std::memcpy(&fp, fp_, sizeof(fp));
    scene_host s{ make_ref<tri_t>(prims, n_prims, nodes, n_nodes, indices, n_idx) };
    fill_scene(s, fp, normals16, n_prims, (mat_pod const*)mats_, (light_pod const*)lights_);
    bool use_matte = fp.n_materials > 0 && ((mat_pod const*)mats_)[0].kind == 1;

    vec4 bg(fp.bg[0], fp.bg[1], fp.bg[2], fp.bg[3]);
    vec4 amb(fp.ambient[0], fp.ambient[1], fp.ambient[2], fp.ambient[3]);

    auto lb = s.lights.data();
    auto le = s.lights.data() + s.lights.size();
    tri_ref_t const* pb = &s.ref;
    tri_ref_t const* pe = &s.ref + 1;

#define VSB_DISPATCH(KERNEL_EXPR, IS_RNG)                                                          \
    do {                                                                                           \
        auto k__ = KERNEL_EXPR;                                                                    \
        if (mode == 0) { scalar_frame(k__, fp, rgba, nthreads, basis_out); return 0.0; }           \
    } while (0)

    if (fp.kernel == 2)
    {
        if (mode == 0)
        {
            ao_kernel<ray_t> k{ pb, pe, s.normals.data(), bg, fp.ao_samples, fp.ao_radius };
            scalar_frame(k, fp, rgba, nthreads, basis_out);
            return 0.0;
        }
#if VSNRAY_SIMD_ISA_GE(VSNRAY_SIMD_ISA_AVX)
        if (mode == 1)
        {
            ao_tiled_adapter<simd::ray8> k{ { pb, pe, s.normals.data(), bg, fp.ao_samples, fp.ao_radius } };
            return tiled_frame<simd::ray8>(k, pixel_sampler::uniform_type{}, fp, rgba, nthreads, reps, warmup);
        }
#endif
        ao_tiled_adapter<ray_t> k{ { pb, pe, s.normals.data(), bg, fp.ao_samples, fp.ao_radius } };
        return tiled_frame<ray_t>(k, pixel_sampler::uniform_type{}, fp, rgba, nthreads, reps, warmup);
    }

    // built-in kernels: plastic or matte material list
    auto run = [&](auto kparams) -> double
    {
        using KP = decltype(kparams);
        if (fp.kernel == 0)
        {
            simple::kernel<KP> k{ kparams };
            if (mode == 0) { scalar_frame(no_sampler_adapter<simple::kernel<KP>>{ k }, fp, rgba, nthreads, basis_out); return 0.0; }
#if VSNRAY_SIMD_ISA_GE(VSNRAY_SIMD_ISA_AVX)
            if (mode == 1) return tiled_frame_ps<simd::ray8>(k, fp, rgba, nthreads, reps, warmup);
#endif
            return tiled_frame_ps<ray_t>(k, fp, rgba, nthreads, reps, warmup);
        }
        else if (fp.kernel == 1)
        {
            whitted::kernel<KP> k{ kparams };
            if (mode == 0) { scalar_frame(no_sampler_adapter<whitted::kernel<KP>>{ k }, fp, rgba, nthreads, basis_out); return 0.0; }
#if VSNRAY_SIMD_ISA_GE(VSNRAY_SIMD_ISA_AVX)
            if (mode == 1) return tiled_frame_ps<simd::ray8>(k, fp, rgba, nthreads, reps, warmup);
#endif
            return tiled_frame_ps<ray_t>(k, fp, rgba, nthreads, reps, warmup);
        }
        else
        {
            pathtracing::kernel<KP> k{ kparams };
            if (mode == 0) { scalar_frame(k, fp, rgba, nthreads, basis_out); return 0.0; }
#if VSNRAY_SIMD_ISA_GE(VSNRAY_SIMD_ISA_AVX)
            if (mode == 1) return tiled_frame<simd::ray8>(k, pixel_sampler::jittered_blend_type{}, fp, rgba, nthreads, reps, warmup);
#endif
            return tiled_frame<ray_t>(k, pixel_sampler::jittered_blend_type{}, fp, rgba, nthreads, reps, warmup);
        }
    };
#undef VSB_DISPATCH

    if (use_matte)
    {
        return run(make_kernel_params(normals_per_face_binding{}, pb, pe, s.normals.data(),
                                      s.mattes.data(), lb, le, fp.num_bounces, fp.epsilon, bg, amb));
    }
    return run(make_kernel_params(normals_per_face_binding{}, pb, pe, s.normals.data(),
                                  s.plastics.data(), lb, le, fp.num_bounces, fp.epsilon, bg, amb));
}

// scalar parity driver with sched_params-style view / projection matrices (column-major 16 floats)
// and an optional PF_DEPTH32F target; the inverses are the reference's inverse() like the
// schedulers compute them (detail/tiled_sched.inl:241-242).  inv_out (32 floats) may be NULL.
void ref_frame_matrix(void const* fp_, float const* view16, float const* proj16, void const* prims, size_t n_prims,
                      void const* nodes, size_t n_nodes, uint32_t const* indices, size_t n_idx, float const* normals16,
                      void const* mats_, void const* lights_, float* rgba, float* depth, int nthreads, float* inv_out)
{
    g_mcam.on = true;
    g_mcam.view = mat4(view16); g_mcam.proj = mat4(proj16);
    g_mcam.inv_view = inverse(g_mcam.view); g_mcam.inv_proj = inverse(g_mcam.proj);
    g_mcam.depth = depth;
    if (inv_out)
    {
        std::memcpy(inv_out, g_mcam.inv_view.data(), 64);
        std::memcpy(inv_out + 16, g_mcam.inv_proj.data(), 64);
    }
    ref_frame(fp_, prims, n_prims, nodes, n_nodes, indices, n_idx, normals16, mats_, lights_, rgba, nthreads, 0, 1, 0, nullptr);
    g_mcam = matrix_cam();
}

// the reference camera's matrices for a frame_params view (detail/camera.inl:10-65), 2 x 16 floats
void ref_camera_matrices(void const* fp_, float z_near, float z_far, float* view16, float* proj16)
{
    frame_params fp;
    std::memcpy(&fp, fp_, sizeof(fp));
    camera cam;
    cam.look_at(vec3(fp.eye[0], fp.eye[1], fp.eye[2]), vec3(fp.center[0], fp.center[1], fp.center[2]), vec3(fp.up[0], fp.up[1], fp.up[2]));
    cam.perspective(fp.fovy, fp.aspect, z_near, z_far);
    std::memcpy(view16, cam.get_view_matrix().data(), 64);
    std::memcpy(proj16, cam.get_proj_matrix().data(), 64);
}

// Timed ray-list baseline: simd::ray8 packets of 8 consecutive rays, all threads.
// Returns best seconds over reps.  hits_out may be NULL.
double ref_query_timed(int any, void const* prims, size_t n_prims, void const* nodes, size_t n_nodes,
                       uint32_t const* indices, size_t n_idx, void const* rays, size_t n_rays,
                       float max_t, int nthreads, int reps, int warmup, uint64_t* n_hits_out)
{
    auto ref = make_ref<tri_t>(prims, n_prims, nodes, n_nodes, indices, n_idx);
    auto r = reinterpret_cast<ray_t const*>(rays);
    double best = 1e30;
    std::atomic<uint64_t> hits(0);
    for (int it = 0; it < warmup + reps; ++it)
    {
        hits = 0;
        auto t0 = std::chrono::steady_clock::now();
#if VSNRAY_SIMD_ISA_GE(VSNRAY_SIMD_ISA_AVX)
        size_t n_packets = n_rays / 8;
        parallel_for(n_packets, nthreads, [&](size_t b, size_t e)
        {
            uint64_t local = 0;
            for (size_t p = b; p < e; ++p)
            {
                std::array<ray_t, 8> rs;
                for (int k = 0; k < 8; ++k) rs[k] = r[p * 8 + k];
                simd::ray8 r8 = simd::pack(rs);
                default_intersector isect;
                if (any)
                {
                    auto hr = any_hit(r8, &ref, &ref + 1, simd::float8(max_t), isect);
                    local += (uint64_t)__builtin_popcount((unsigned)_mm256_movemask_ps(hr.hit.f));
                }
                else
                {
                    auto hr = closest_hit(r8, &ref, &ref + 1, isect);
                    local += (uint64_t)__builtin_popcount((unsigned)_mm256_movemask_ps(hr.hit.f));
                }
            }
            hits += local;
        });
#else
        parallel_for(n_rays, nthreads, [&](size_t b, size_t e)
        {
            uint64_t local = 0;
            for (size_t i = b; i < e; ++i)
            {
                default_intersector isect;
                if (any) local += any_hit(r[i], &ref, &ref + 1, max_t, isect).hit ? 1 : 0;
                else     local += closest_hit(r[i], &ref, &ref + 1, isect).hit ? 1 : 0;
            }
            hits += local;
        });
#endif
        auto t1 = std::chrono::steady_clock::now();
        double dt = std::chrono::duration<double>(t1 - t0).count();
        if (it >= warmup && dt < best) best = dt;
    }
    if (n_hits_out) *n_hits_out = hits.load();
    return best;
}

int ref_simd_width()
{
#if VSNRAY_SIMD_ISA_GE(VSNRAY_SIMD_ISA_AVX)
    return 8;
#else
    return 1;
#endif
}

} // extern "C"
