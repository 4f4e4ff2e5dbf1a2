// dropin_example.cpp -- a reference-style program written against the host adaptor headers
// (include/visionaray_b200/), i.e. what a user of tu500/visionaray's cuda_sched path writes after
// switching.  Used by tests/test_dropin_cpp.py: it dumps hit records and frames
// for the same seeded scene the Python tests regenerate and check against the oracle.
//
//   dropin_example <out_dir>      exit 0 ok, 2 known-answer failure, 3 no CUDA device / library error
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <vector>

#include <visionaray_b200/scheduler.h>
#include <visionaray_b200/obj_loader.h>
#include "../../visionaray_b200/csrc/host/scenes.h"

namespace visionaray = visionaray_b200;      // the one-line switch
using namespace visionaray;

template <typename T>
static void dump(std::string const& path, T const* p, size_t n)
{
    std::ofstream f(path, std::ios::binary);
    f.write(reinterpret_cast<char const*>(p), std::streamsize(n * sizeof(T)));
}

int main(int argc, char** argv)
{
    std::string out = argc > 1 ? argv[1] : ".";
    try
    {
        using triangle_type = basic_triangle<3, float>;

        // --- GetNormal.BVH (test/unittests/get_normal.cpp:17-46) ---
        {
            std::vector<triangle_type> t(2);
            t[0] = scenes::make_tri(vec3(-1, -1, 1), vec3(1, -1, 1), vec3(1, 1, 1), 0, 0);
            t[1] = scenes::make_tri(vec3(1, -1, -1), vec3(-1, -1, -1), vec3(-1, 1, -1), 1, 1);
            auto bvh = build<index_bvh<triangle_type>>(t.data(), t.size());
            b200_index_bvh<triangle_type> d_bvh(bvh);
            b200_bvh_ref refs[1] = { d_bvh.ref() };
            ray r(vec3(0.5f, -0.5f, 2.0f), vec3(0.0f, 0.0f, -1.0f));
            auto hr = closest_hit(r, refs, refs + 1);
            if (!hr.hit || hr.t != 1.0f || hr.prim_id != 0) { std::cerr << "known-answer test failed\n"; return 2; }
            if (any_hit(r, refs, refs + 1, 0.5f).hit) { std::cerr << "any_hit max_t failed\n"; return 2; }
        }

        // --- generic_primitive<triangle, sphere>: one tree, both kinds (src/examples/generic_primitive) ---
        {
            using generic_type = generic_primitive<basic_triangle<3, float>, basic_sphere<float>>;
            triangle_type tri = scenes::make_tri(vec3(-1, -1, 1), vec3(1, -1, 1), vec3(1, 1, 1), 0, 0);
            basic_sphere<float> sph(vec3(0.0f, 0.0f, -2.0f), 1.0f);
            sph.prim_id = 1; sph.geom_id = 1;
            std::vector<generic_type> gp = { generic_type(tri), generic_type(sph) };
            auto gbvh = build<index_bvh<generic_type>>(gp.data(), gp.size());
            b200_index_bvh<generic_type> d_gbvh(gbvh);
            b200_bvh_ref grefs[1] = { d_gbvh.ref() };
            auto h0 = closest_hit(ray(vec3(0.5f, -0.5f, 5.0f), vec3(0.0f, 0.0f, -1.0f)), grefs, grefs + 1);
            auto h1 = closest_hit(ray(vec3(-0.5f, 0.5f, 5.0f), vec3(0.0f, 0.0f, -1.0f)), grefs, grefs + 1);
            if (!h0.hit || h0.prim_id != 0 || h0.t != 4.0f) { std::cerr << "generic primitive: triangle known answer failed\n"; return 2; }
            if (!h1.hit || h1.prim_id != 1 || !(h1.t > 6.29f && h1.t < 6.30f)) { std::cerr << "generic primitive: sphere known answer failed\n"; return 2; }
        }

        // --- scene: UV sphere, binned-SAH index_bvh on the host, upload + GPU re-encode ---
        std::vector<triangle_type> tris;
        scenes::uv_sphere(tris, 50, 50, 1.0f);
        auto bvh = build<index_bvh<triangle_type>>(tris.data(), tris.size());
        std::vector<vec3> normals;
        scenes::face_normals(normals, tris);
        b200_index_bvh<triangle_type> d_bvh(bvh);
        d_bvh.set_normals(normals.data(), normals.size());
        b200_bvh_ref prims[1] = { d_bvh.ref() };

        // --- ray list: closest_hit / any_hit ---
        std::vector<ray> rays(20000);
        scenes::random_rays(rays.data(), rays.size(), 3, -1.5f, 1.5f);
        std::vector<hit_record> wide(rays.size()), exact(rays.size()), occ(rays.size());
        closest_hit(rays.data(), rays.size(), prims, prims + 1, wide.data());
        closest_hit(rays.data(), rays.size(), prims, prims + 1, exact.data(), true);
        any_hit(rays.data(), rays.size(), prims, prims + 1, 0.5f, occ.data());
        dump(out + "/hits_wide.bin", wide.data(), wide.size());
        dump(out + "/hits_exact.bin", exact.data(), exact.size());
        dump(out + "/hits_any.bin", occ.data(), occ.size());

        // --- multi_hit<16>: a ray through the centre of the sphere crosses the mesh twice ---
        {
            ray through(vec3(0.013f, 0.021f, 3.0f), vec3(0.0f, 0.0f, -1.0f));
            auto mh = multi_hit<16>(through, prims, prims + 1);
            if (!mh[0].hit || !mh[1].hit || mh[2].hit || !(mh[0].t < mh[1].t)) { std::cerr << "multi_hit known answer failed\n"; return 2; }
            if (mh[0].t != closest_hit(through, prims, prims + 1).t) { std::cerr << "multi_hit[0] != closest_hit\n"; return 2; }
            std::vector<hit_record> all(rays.size() * 4);
            multi_hit<4>(rays.data(), rays.size(), prims, prims + 1, all.data(), true);
            dump(out + "/hits_multi4.bin", all.data(), all.size());
        }

        // --- the same mesh with the BVH built on the GPU (replaces the host build + upload) ---
        {
            auto g_bvh = build<b200_index_bvh<triangle_type>>(tris.data(), tris.size());
            b200_bvh_ref gp[1] = { g_bvh.ref() };
            std::vector<hit_record> built(rays.size());
            closest_hit(rays.data(), rays.size(), gp, gp + 1, built.data());
            dump(out + "/hits_gpu_built.bin", built.data(), built.size());
            index_bvh<triangle_type> back = g_bvh.download(tris.data());      // an ordinary host index_bvh again
            if (back.num_nodes() != g_bvh.num_nodes() || back.num_indices() != tris.size()) { std::cerr << "download size\n"; return 2; }
            b200_index_bvh<triangle_type> again(back);                        // ... that uploads like a host-built one
            b200_bvh_ref ap[1] = { again.ref() };
            std::vector<hit_record> h1(rays.size()), h2(rays.size());
            closest_hit(rays.data(), rays.size(), gp, gp + 1, h1.data(), true);
            closest_hit(rays.data(), rays.size(), ap, ap + 1, h2.data(), true);
            if (std::memcmp(h1.data(), h2.data(), h1.size() * sizeof(hit_record)) != 0) { std::cerr << "downloaded tree != device tree\n"; return 2; }
        }

        // --- frame(): Whitted kernel, one point light at the eye, plastic material ---
        int w = 160, h = 120;
        pinhole_camera cam;
        cam.look_at(vec3(0.0f, 0.0f, 3.0f), vec3(0.0f), vec3(0.0f, 1.0f, 0.0f));
        cam.perspective(45.0f * 3.14159265358979323846f / 180.0f, float(w) / float(h), 0.001f, 1000.0f);
        cam.set_viewport(0, 0, w, h);

        std::vector<plastic<float>> materials(1);
        materials[0].set_ca(vec3(0.2f)); materials[0].set_ka(1.0f);
        materials[0].set_cd(vec3(0.8f)); materials[0].set_kd(1.0f);
        materials[0].set_cs(vec3(0.1f)); materials[0].set_ks(1.0f);
        materials[0].set_specular_exp(32.0f);
        std::vector<point_light<float>> lights(1);
        lights[0].set_position(vec3(0.0f, 0.0f, 3.0f));

        auto kparams = make_kernel_params(normals_per_face_binding{}, prims, prims + 1, (vec3 const*)nullptr, materials,
                                          lights.data(), lights.data() + lights.size(), 4, 1e-3f,
                                          vec4(0.1f, 0.4f, 1.0f, 1.0f), vec4(0.0f));
        whitted::kernel<decltype(kparams)> kernel{ kparams };

        simple_buffer_rt<PF_RGBA32F, PF_UNSPECIFIED> host_rt;
        host_rt.resize(w, h);
        b200_sched<ray> sched;
        auto sparams = make_sched_params(pixel_sampler::uniform_type{}, cam, host_rt);
        sched.frame(kernel, sparams);
        dump(out + "/frame_whitted.bin", host_rt.color(), size_t(w) * h);

        gpu_buffer_rt<PF_RGBA32F, PF_UNSPECIFIED> dev_rt;
        dev_rt.resize(w, h);
        auto sparams_dev = make_sched_params(cam, dev_rt);
        sched.frame(kernel, sparams_dev);
        std::vector<vec4> back(size_t(w) * h);
        dev_rt.download(back.data());
        if (std::memcmp(back.data(), host_rt.color(), back.size() * sizeof(vec4)) != 0) { std::cerr << "gpu_buffer_rt != simple_buffer_rt\n"; return 2; }

        // --- progressive path tracing: jittered_blend, frame_num = 1..3 ---
        std::vector<matte<float>> mattes(1);
        mattes[0].set_ca(vec3(0.2f)); mattes[0].set_ka(1.0f); mattes[0].set_cd(vec3(0.8f)); mattes[0].set_kd(1.0f);
        auto kp2 = make_kernel_params(normals_per_face_binding{}, prims, prims + 1, (vec3 const*)nullptr, mattes,
                                      lights.data(), lights.data() + lights.size(), 4, 1e-3f,
                                      vec4(0.1f, 0.4f, 1.0f, 1.0f), vec4(1.0f));
        pathtracing::kernel<decltype(kp2)> pt{ kp2 };
        host_rt.clear_color_buffer();
        auto sp2 = make_sched_params(pixel_sampler::jittered_blend_type{}, cam, host_rt);
        for (unsigned frame_num = 1; frame_num <= 3; ++frame_num) sched.frame(pt, sp2, frame_num);
        dump(out + "/frame_pt.bin", host_rt.color(), size_t(w) * h);

        // --- sched_params from the camera's matrices + a depth buffer (the form GL compositing uses):
        //     SSAA 4x Whitted into RGBA32F + DEPTH32F, host and device targets ---
        pinhole_camera cam2;
        cam2.look_at(vec3(0.0f, 0.0f, 3.0f), vec3(0.0f), vec3(0.0f, 1.0f, 0.0f));
        cam2.perspective(45.0f * 3.14159265358979323846f / 180.0f, float(w) / float(h), 0.1f, 10.0f);
        simple_buffer_rt<PF_RGBA32F, PF_DEPTH32F> depth_rt;
        depth_rt.resize(w, h);
        auto sp3 = make_sched_params(pixel_sampler::ssaa_type<4>{}, cam2.get_view_matrix(), cam2.get_proj_matrix(), depth_rt);
        sched.frame(kernel, sp3);
        dump(out + "/frame_matrix_color.bin", depth_rt.color(), size_t(w) * h);
        dump(out + "/frame_matrix_depth.bin", depth_rt.depth(), size_t(w) * h);
        gpu_buffer_rt<PF_RGBA32F, PF_DEPTH32F> dev_depth_rt;
        dev_depth_rt.resize(w, h);
        auto sp4 = make_sched_params(pixel_sampler::ssaa_type<4>{}, cam2.get_view_matrix(), cam2.get_proj_matrix(), dev_depth_rt);
        sched.frame(kernel, sp4);
        std::vector<float> zback(size_t(w) * h);
        dev_depth_rt.download_depth(zback.data());
        if (std::memcmp(zback.data(), depth_rt.depth(), zback.size() * sizeof(float)) != 0) { std::cerr << "gpu depth != host depth\n"; return 2; }

        // --- a sphere bvh rendered with kernel_params that carry no normals (get_normal(hr, sphere)) ---
        {
            std::vector<basic_sphere<float>> balls(64);
            for (size_t i = 0; i < balls.size(); ++i)
            {
                balls[i] = basic_sphere<float>(vec3(-1.4f + 0.4f * float(i % 8), -1.4f + 0.4f * float(i / 8), 0.1f * float(i % 3)), 0.17f);
                balls[i].prim_id = unsigned(i); balls[i].geom_id = 0;
            }
            auto sb = build<index_bvh<basic_sphere<float>>>(balls.data(), balls.size());
            b200_index_bvh<basic_sphere<float>> d_sb(sb);
            b200_bvh_ref sp[1] = { d_sb.ref() };
            auto skp = make_kernel_params(sp, sp + 1, materials, lights.data(), lights.data() + lights.size(), 4, 1e-3f,
                                          vec4(0.1f, 0.4f, 1.0f, 1.0f), vec4(0.0f));
            whitted::kernel<decltype(skp)> sk{ skp };
            b200_sched<ray> s2;
            s2.set_reference_order(true);
            s2.frame(sk, sparams);
            dump(out + "/frame_spheres.bin", host_rt.color(), size_t(w) * h);
            dump(out + "/spheres.bin", balls.data(), balls.size());
        }

        // --- a spot_light list instead of point lights (spot_light.h) ---
        {
            std::vector<spot_light<float>> spots(1);
            spots[0].set_position(vec3(0.5f, 1.0f, 3.0f));
            vec3 dir = normalize(vec3(0.0f) - vec3(0.5f, 1.0f, 3.0f));
            spots[0].set_spot_direction(dir);
            spots[0].set_spot_cutoff(0.3f);
            spots[0].set_spot_exponent(6.0f);
            auto kp_s = make_kernel_params(normals_per_face_binding{}, prims, prims + 1, (vec3 const*)nullptr, materials,
                                           spots.data(), spots.data() + spots.size(), 4, 1e-3f,
                                           vec4(0.1f, 0.4f, 1.0f, 1.0f), vec4(0.0f));
            whitted::kernel<decltype(kp_s)> ks{ kp_s };
            b200_sched<ray> s3;
            s3.set_reference_order(true);
            s3.frame(ks, sparams);
            dump(out + "/frame_spot.bin", host_rt.color(), size_t(w) * h);
            vsb_point_light fl = spots[0].flatten();
            dump(out + "/spot_light.bin", &fl, 1);
        }

        // --- other colour formats: the same Whitted frame into PF_RGBA8 (host) and PF_RGB8 (device) targets ---
        {
            simple_buffer_rt<PF_RGBA8, PF_UNSPECIFIED> rt8;
            rt8.resize(w, h);
            auto sp8 = make_sched_params(pixel_sampler::uniform_type{}, cam, rt8);
            b200_sched<ray> s8;
            s8.set_reference_order(true);
            s8.frame(kernel, sp8);
            dump(out + "/frame_whitted_rgba8.bin", rt8.color(), size_t(w) * h);
            gpu_buffer_rt<PF_RGB8, PF_UNSPECIFIED> rt3;
            rt3.resize(w, h);
            auto sp3b = make_sched_params(pixel_sampler::ssaa_type<2>{}, cam, rt3);
            s8.frame(kernel, sp3b);
            std::vector<unorm8x3> px(size_t(w) * h);
            rt3.download(px.data());
            dump(out + "/frame_whitted_ssaa2_rgb8.bin", px.data(), px.size());
        }

        // --- kernel_params with a surface description (kernels.h:395-509): normals per vertex, colours per
        //     vertex, tex coords + one texture_ref per geom_id, generic_material list; Whitted, host target.
        //     The arrays are dumped so that the test can hand the same ones to the oracle. ---
        {
            using material_type = generic_material<emissive<float>, matte<float>, mirror<float>, plastic<float>>;
            std::vector<triangle_type> st = tris;
            for (size_t i = 0; i < st.size(); ++i) st[i].geom_id = unsigned((i / 3) % 4);
            auto sbvh = build<index_bvh<triangle_type>>(st.data(), st.size());
            b200_index_bvh<triangle_type> d_sbvh(sbvh);
            b200_bvh_ref sprims[1] = { d_sbvh.ref() };

            std::vector<vec3> vn(st.size() * 3), vc(st.size() * 3);
            std::vector<vec2> tc(st.size() * 3);
            for (size_t i = 0; i < st.size(); ++i)
            {
                vec3 v[3] = { st[i].v1, st[i].v1 + st[i].e1, st[i].v1 + st[i].e2 };
                for (int k = 0; k < 3; ++k)
                {
                    vn[i * 3 + k] = normalize(v[k]);                                  // the sphere's smooth normal
                    vc[i * 3 + k] = vec3(0.5f + 0.5f * v[k].x, 0.5f + 0.5f * v[k].y, 0.5f + 0.5f * v[k].z);
                    tc[i * 3 + k].x = 1.7f * v[k].x + 0.4f * v[k].y; tc[i * 3 + k].y = 1.3f * v[k].z - 0.2f;
                }
            }
            int tw = 24, th = 16;
            std::vector<unorm8x4> texels(size_t(tw) * th);
            for (int y = 0; y < th; ++y)
                for (int x = 0; x < tw; ++x)
                {
                    unsigned char c = (unsigned char)(((x / 3 + y / 3) & 1) ? 230 : 70);
                    texels[size_t(y) * tw + x] = unorm8x4{ c, (unsigned char)(c / 2 + x * 4), (unsigned char)(255 - c + y), 255 };
                }
            std::vector<texture_ref2d> textures(4);             // geom_id 2 and 3 stay without a texture
            textures[0] = texture_ref2d(tw, th); textures[0].reset(texels.data());
            textures[0].set_filter_mode(Linear); textures[0].set_address_mode(0, Wrap); textures[0].set_address_mode(1, Mirror);
            textures[1] = texture_ref2d(tw, th); textures[1].reset(texels.data());
            textures[1].set_filter_mode(Nearest); textures[1].set_address_mode(Clamp);

            plastic<float> pl; pl.set_ca(vec3(0.2f)); pl.set_ka(1.0f); pl.set_cd(vec3(0.8f)); pl.set_kd(1.0f); pl.set_cs(vec3(0.1f)); pl.set_ks(1.0f); pl.set_specular_exp(32.0f);
            matte<float> ma; ma.set_ca(vec3(0.2f)); ma.set_ka(1.0f); ma.set_cd(vec3(0.7f, 0.5f, 0.3f)); ma.set_kd(1.0f);
            mirror<float> mi; mi.set_cr(vec3(0.9f, 0.85f, 0.8f)); mi.set_kr(0.8f); mi.set_ior(vec3(0.2f, 0.9f, 1.1f)); mi.set_absorption(vec3(3.9f, 2.4f, 2.2f));
            emissive<float> em; em.set_ce(vec3(1.0f, 0.9f, 0.7f)); em.set_ls(2.0f);
            std::vector<material_type> gmats = { material_type(pl), material_type(ma), material_type(mi), material_type(em) };
            std::vector<vsb_material> flat;
            for (auto const& m : gmats) flat.push_back(m.flatten());

            auto skp = make_kernel_params(normals_per_vertex_binding{}, colors_per_vertex_binding{}, sprims, sprims + 1, vn.data(),
                                          tc.data(), gmats, vc.data(), textures, lights.data(), lights.data() + lights.size(),
                                          4, 1e-3f, vec4(0.1f, 0.4f, 1.0f, 1.0f), vec4(0.3f, 0.3f, 0.3f, 1.0f));
            whitted::kernel<decltype(skp)> sk{ skp };
            b200_sched<ray> ssched;
            ssched.set_reference_order(true);
            ssched.frame(sk, sparams);
            dump(out + "/surface_whitted.bin", host_rt.color(), size_t(w) * h);
            // attached once to the device tree, frames pass nullptr: same pixels
            d_sbvh.set_surface(skp.surface);
            auto skp2 = skp; skp2.surface.normals = nullptr;
            whitted::kernel<decltype(skp2)> sk2{ skp2 };
            std::vector<vec4> first(host_rt.color(), host_rt.color() + size_t(w) * h);
            ssched.frame(sk2, sparams);
            if (std::memcmp(first.data(), host_rt.color(), first.size() * sizeof(vec4)) != 0) { std::cerr << "attached surface != per-frame surface\n"; return 2; }
            dump(out + "/surface_normals.bin", vn.data(), vn.size());
            dump(out + "/surface_colors.bin", vc.data(), vc.size());
            dump(out + "/surface_tex_coords.bin", tc.data(), tc.size());
            dump(out + "/surface_texels.bin", texels.data(), texels.size());
            dump(out + "/surface_materials.bin", flat.data(), flat.size());
        }

        // --- a scene from an OBJ file the way the reference's viewer sets it up (src/viewer/viewer.cpp:480-541):
        //     load_obj -> build<index_bvh> -> make_kernel_params(normals_per_vertex_binding, ..., shading_normals,
        //     tex_coords, materials, textures, ...) -> whitted kernel -> frame().  argv[2] = the .obj file. ---
        if (argc > 2)
        {
            model mod;
            load_obj(argv[2], mod);
            if (mod.shading_normals.size() != 3 * mod.primitives.size()) { std::cerr << "obj: the example needs normals on every face\n"; return 2; }
            auto obvh = build<index_bvh<triangle_type>>(mod.primitives.data(), mod.primitives.size());
            b200_index_bvh<triangle_type> d_obvh(obvh);
            b200_bvh_ref oprims[1] = { d_obvh.ref() };
            std::vector<point_light<float>> olights(1);
            olights[0].set_position(vec3(2.5f, 4.0f, 3.0f));
            auto okp = make_kernel_params(normals_per_vertex_binding{}, oprims, oprims + 1, mod.shading_normals.data(), mod.tex_coords.data(),
                                          mod.materials, mod.textures, olights.data(), olights.data() + olights.size(), 3, 1e-3f,
                                          vec4(0.1f, 0.4f, 1.0f, 1.0f), vec4(0.2f, 0.2f, 0.2f, 1.0f));
            whitted::kernel<decltype(okp)> ok{ okp };
            pinhole_camera ocam;
            ocam.look_at(vec3(1.6f, 2.1f, 2.9f), vec3(0.0f, 0.6f, 0.0f));
            ocam.perspective(45.0f * 3.14159265358979323846f / 180.0f, float(224) / float(160), 0.001f, 1000.0f);
            simple_buffer_rt<PF_RGBA32F, PF_UNSPECIFIED> ort;
            ort.resize(224, 160);
            auto osp = make_sched_params(pixel_sampler::uniform_type{}, ocam, ort);
            b200_sched<ray> osched;
            osched.set_reference_order(true);
            osched.frame(ok, osp);
            osched.frame(ok, osp);          // second frame: the surface description is not uploaded again (same pixels)
            dump(out + "/frame_obj.bin", ort.color(), size_t(224) * 160);
        }

        std::printf("dropin_example ok: %zu triangles, %zu nodes, %zu rays\n", tris.size(), bvh.num_nodes(), rays.size());
        return 0;
    }
    catch (std::exception const& e)
    {
        std::cerr << "error: " << e.what() << "\n";
        return 3;
    }
}
