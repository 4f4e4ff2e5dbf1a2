// ref_cuda_harness.cu -- TEST / BASELINE INFRASTRUCTURE ONLY (see oracle/README.md).
//
// The reference's OWN CUDA path for frame(): cuda_sched<basic_ray<float>>::frame() (detail/cuda_sched.inl:60-320) with the
// stock whitted::kernel over a cuda_index_bvh (bvh.h:443-448), compiled from the UNMODIFIED headers under
// /root/reference/include for sm_100a.  It is the "naive GPU" baseline of bench.py (--impl reference-cuda): the same
// operation by the reference's one-thread-per-pixel BVH2 kernel on the same B200.  Nothing of this repository's kernels
// is used.  gpu_buffer_rt.h is not usable here (it includes cpu_buffer_rt.h -> gl/compositing.h -> the generated
// visionaray/config.h and GL), so the render target is a harness-local struct with the duck-typed contract of
// render_target.h:45-101 over a thrust::device_vector.
#include <cstdint>
#include <cstring>
#include <vector>

#include <cuda_runtime.h>
#include <thrust/device_vector.h>

#include <visionaray/math/math.h>
#include <visionaray/bvh.h>
#include <visionaray/camera.h>
#include <visionaray/kernels.h>
#include <visionaray/material.h>
#include <visionaray/point_light.h>
#include <visionaray/render_target.h>
#include <visionaray/scheduler.h>

using namespace visionaray;

namespace
{

struct device_rt
{
    using color_type = vector<4, float>;
    using depth_type = typename pixel_traits<PF_UNSPECIFIED>::type;
    using ref_type = render_target_ref<PF_RGBA32F, PF_UNSPECIFIED>;

    thrust::device_vector<color_type> color_buffer;
    int w = 0, h = 0;

    void resize(int width, int height) { w = width; h = height; color_buffer.resize(size_t(w) * h); }
    int width() const { return w; }
    int height() const { return h; }
    color_type* color() { return thrust::raw_pointer_cast(color_buffer.data()); }
    depth_type* depth() { return nullptr; }
    ref_type ref() { return ref_type(color(), depth(), w, h); }
    void begin_frame() {}
    void end_frame() {}
};

using triangle_type = basic_triangle<3, float>;
using host_bvh = index_bvh<triangle_type>;
using device_bvh = cuda_index_bvh<triangle_type>;

} // namespace

extern "C"
{

// Renders `reps` Whitted frames of (prims, the reference's own binned-SAH tree) on the GPU; returns the average
// milliseconds per frame (CUDA events), the last frame in rgba_out (host, w * h * 4 floats; may be NULL).
// eye / center / up, fovy in radians; one point light at `light`; plastic default material like the CPU harness.
float ref_cuda_whitted(const void* tris64, size_t n_tris, int w, int h, const float* eye, const float* center, const float* up,
                       float fovy, const float* light, unsigned num_bounces, float epsilon, const float* bg4,
                       int reps, int warmup, float* rgba_out, float* build_upload_ms)
{
    const triangle_type* tris = reinterpret_cast<const triangle_type*>(tris64);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    host_bvh bvh = build<host_bvh>(tris, n_tris);
    device_bvh d_bvh(bvh);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    if (build_upload_ms) cudaEventElapsedTime(build_upload_ms, e0, e1);

    std::vector<vector<3, float>> normals(n_tris);
    for (size_t i = 0; i < n_tris; ++i) normals[i] = normalize(cross(tris[i].e1, tris[i].e2));
    thrust::device_vector<vector<3, float>> d_normals(normals);

    plastic<float> m;
    m.set_ca(from_rgb(0.2f, 0.2f, 0.2f)); m.set_cd(from_rgb(0.8f, 0.8f, 0.8f)); m.set_cs(from_rgb(0.1f, 0.1f, 0.1f));
    m.set_ka(1.0f); m.set_kd(1.0f); m.set_ks(1.0f); m.set_specular_exp(32.0f);
    thrust::device_vector<plastic<float>> d_materials(1, m);

    point_light<float> l;
    l.set_cl(vec3(1.0f, 1.0f, 1.0f)); l.set_kl(1.0f);
    l.set_position(vec3(light[0], light[1], light[2]));
    l.set_constant_attenuation(1.0f); l.set_linear_attenuation(0.0f); l.set_quadratic_attenuation(0.0f);
    thrust::device_vector<point_light<float>> d_lights(1, l);

    thrust::device_vector<typename device_bvh::bvh_ref> d_refs(1, d_bvh.ref());

    auto kparams = make_kernel_params(
        normals_per_face_binding{},
        thrust::raw_pointer_cast(d_refs.data()), thrust::raw_pointer_cast(d_refs.data()) + 1,
        thrust::raw_pointer_cast(d_normals.data()),
        thrust::raw_pointer_cast(d_materials.data()),
        thrust::raw_pointer_cast(d_lights.data()), thrust::raw_pointer_cast(d_lights.data()) + 1,
        num_bounces, epsilon, vec4(bg4[0], bg4[1], bg4[2], bg4[3]), vec4(0.0f));
    whitted::kernel<decltype(kparams)> kernel;
    kernel.params = kparams;

    camera cam;
    cam.perspective(fovy, float(w) / float(h), 0.001f, 1000.0f);
    cam.look_at(vec3(eye[0], eye[1], eye[2]), vec3(center[0], center[1], center[2]), vec3(up[0], up[1], up[2]));

    device_rt rt;
    rt.resize(w, h);
    auto sparams = make_sched_params(pixel_sampler::uniform_type{}, cam, rt);
    cuda_sched<basic_ray<float>> sched;
    for (int i = 0; i < warmup; ++i) sched.frame(kernel, sparams);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    for (int i = 0; i < reps; ++i) sched.frame(kernel, sparams);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, e0, e1);
    if (rgba_out) cudaMemcpy(rgba_out, rt.color(), size_t(w) * h * 16, cudaMemcpyDeviceToHost);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return cudaGetLastError() == cudaSuccess ? ms / float(reps > 0 ? reps : 1) : -1.0f;
}

} // extern "C"
