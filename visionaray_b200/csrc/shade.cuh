// shade.cuh -- device restatement of the frame() callers of the ray-query path: primary ray
// generation (detail/sched_common.h:130-150), the built-in kernels simple / whitted /
// pathtracing (detail/{simple,whitted,pathtracing}.inl) with the plastic / matte + point-light
// subset (detail/material/{plastic,matte}.inl, brdf.h, detail/point_light.inl), and the AO
// kernel of src/examples/ao/main.cpp:183-246.
//
// Everything that DEFINES A RAY (primary rays, hit positions, light / reflection / sampled
// directions, offsets) uses single-rounding arithmetic in the reference's operation order, so
// secondary rays are the same rays the reference traces (up to libm-vs-CUDA sinf/cosf/powf
// ulps).  Colour arithmetic is free to use FMAs: pixel parity is a 1/255 tolerance.
#pragma once

#include "vsb_common.cuh"

namespace vsb
{

struct V3 { float x, y, z; };

VSB_HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
VSB_HD V3 vadd(V3 a, V3 b) { return v3(rn_add(a.x, b.x), rn_add(a.y, b.y), rn_add(a.z, b.z)); }
VSB_HD V3 vsub(V3 a, V3 b) { return v3(rn_sub(a.x, b.x), rn_sub(a.y, b.y), rn_sub(a.z, b.z)); }
VSB_HD V3 vmul(V3 a, V3 b) { return v3(rn_mul(a.x, b.x), rn_mul(a.y, b.y), rn_mul(a.z, b.z)); }
VSB_HD V3 vscale(V3 a, float s) { return v3(rn_mul(a.x, s), rn_mul(a.y, s), rn_mul(a.z, s)); }
VSB_HD V3 vneg(V3 a) { return v3(-a.x, -a.y, -a.z); }
VSB_HD float vdot(V3 a, V3 b) { return rn_dot(a.x, a.y, a.z, b.x, b.y, b.z); }
VSB_HD V3 vcross(V3 u, V3 v)
{
    return v3(rn_sub(rn_mul(u.y, v.z), rn_mul(u.z, v.y)), rn_sub(rn_mul(u.z, v.x), rn_mul(u.x, v.z)),
              rn_sub(rn_mul(u.x, v.y), rn_mul(u.y, v.x)));
}
VSB_HD V3 vnormalize(V3 v) { return vscale(v, rn_div(1.0f, rn_sqrt(vdot(v, v)))); }   // v * (1/sqrt(dot))
VSB_HD float vlength(V3 v) { return rn_sqrt(vdot(v, v)); }
VSB_HD V3 vreflect(V3 i, V3 n) { return vsub(vscale(n, rn_mul(2.0f, vdot(n, i))), i); }  // math/detail/vector.inl:685-688
VSB_HD V3 vfaceforward(V3 n, V3 i, V3 nref) { return vdot(nref, i) < 0.0f ? vneg(n) : n; } // :674-681
VSB_HD float max0(float x) { return 0.0f < x ? x : 0.0f; }

#define VSB_PI      3.14159265358979323846f
#define VSB_TWO_PI  6.28318530717958647692f
#define VSB_INV_PI  0.31830988618379067154f

struct Material { float ca[3], ka, cd[3], kd, cs[3], ks, spec_exp; int32_t kind; float pad[2]; };
struct PointLight { float pos[3], kl, cl[3], att_const, att_lin, att_quad, pad[2]; };

struct FrameParams   // == vsb_frame_params
{
    int32_t  kernel, pixel_sampler, width, height;
    float    eye[4], center[4], up[4];
    float    fovy, aspect;
    uint32_t frame_num, seed, num_bounces;
    float    epsilon;
    float    bg[4], ambient[4];
    int32_t  ao_samples;
    float    ao_radius;
    int32_t  n_lights, n_materials;
};

struct FrameCtx
{
    FrameParams p;
    float eye[3], cu[3], cv[3], cw[3];
    const float4*     normals;     // per-face normals, 16 B each, by prim_id
    const Material*   materials;   // by geom_id
    const PointLight* lights;
};

VSB_HD V3 ld3(const float* p) { return v3(p[0], p[1], p[2]); }

// make_primary_ray_impl, pinhole: detail/sched_common.h:130-150
VSB_HD void primary_ray(const FrameCtx& f, float x, float y, V3& ori, V3& dir)
{
    const float u = rn_sub(rn_div(rn_mul(2.0f, rn_add(x, 0.5f)), (float)f.p.width), 1.0f);
    const float v = rn_sub(rn_div(rn_mul(2.0f, rn_add(y, 0.5f)), (float)f.p.height), 1.0f);
    ori = ld3(f.eye);
    dir = vnormalize(vadd(vadd(vscale(ld3(f.cu), u), vscale(ld3(f.cv), v)), ld3(f.cw)));
}

// point_light::intensity -- detail/point_light.inl:11-27
VSB_HD V3 light_intensity(const PointLight& l, V3 pos)
{
    const float dist = vlength(vsub(ld3(l.pos), pos));
    const float att = 1.0f / (l.att_const + l.att_lin * dist + l.att_quad * dist * dist);
    return v3(l.cl[0] * l.kl * att, l.cl[1] * l.kl * att, l.cl[2] * l.kl * att);
}

VSB_HD V3 lambert_f(const Material& m) { return v3(m.cd[0] * m.kd * VSB_INV_PI, m.cd[1] * m.kd * VSB_INV_PI, m.cd[2] * m.kd * VSB_INV_PI); }

// blinn::f -- brdf.h:107-119 (plastic's specular lobe)
VSB_HD V3 blinn_f(const Material& m, V3 n, V3 wo, V3 wi)
{
    const V3 h = vnormalize(vadd(wo, wi));
    const float hdotn = max0(vdot(h, n));
    float sat = vdot(wi, h);
    sat = sat < 0.0f ? 0.0f : (sat > 1.0f ? 1.0f : sat);
    const float sch = powf(1.0f - sat, 5.0f);
    const float nfactor = (m.spec_exp + 2.0f) / (8.0f * VSB_PI);
    const float ph = powf(hdotn, m.spec_exp) * nfactor;
    const float sx = m.cs[0] * m.ks, sy = m.cs[1] * m.ks, sz = m.cs[2] * m.ks;
    return v3((sx + (1.0f - sx) * sch) * ph, (sy + (1.0f - sy) * sch) * ph, (sz + (1.0f - sz) * sch) * ph);
}

// plastic::shade / matte::shade -- detail/material/plastic.inl:18-37, matte.inl:18-39
VSB_HD V3 shade_light(const Material& m, const PointLight& l, V3 pos, V3 n, V3 wo, V3 wi)
{
    const float ndotl = max0(vdot(n, wi));
    V3 f = lambert_f(m);
    if (m.kind == 0) { const V3 s = blinn_f(m, n, wo, wi); f = v3(f.x + s.x, f.y + s.y, f.z + s.z); }
    const V3 li = light_intensity(l, pos);
    return v3(f.x * VSB_PI * li.x * ndotl, f.y * VSB_PI * li.y * ndotl, f.z * VSB_PI * li.z * ndotl);
}

// make_orthonormal_basis -- math/detail/vector3.inl:357-367
VSB_HD void ortho_basis(V3& u, V3& v, V3 w)
{
    if (fabsf(w.x) > fabsf(w.y)) v = vnormalize(v3(-w.z, 0.0f, w.x));
    else                         v = vnormalize(v3(0.0f, w.z, -w.y));
    u = vcross(v, w);
}

// cosine_sample_hemisphere -- sampling.h:61-71
VSB_HD V3 cosine_sample_hemisphere(float u1, float u2)
{
    const float r = rn_sqrt(u1);
    const float theta = rn_mul(VSB_TWO_PI, u2);
    return v3(rn_mul(r, cosf(theta)), rn_mul(r, sinf(theta)), rn_sqrt(max0(rn_sub(1.0f, u1))));
}

VSB_HD V3 hemisphere_dir(V3 u, V3 v, V3 w, V3 sp)
{
    return vnormalize(vadd(vadd(vscale(u, sp.x), vscale(v, sp.y)), vscale(w, sp.z)));
}

// lambertian::sample_f -- brdf.h:43-61.  Draw order: cosine_sample_hemisphere(next(), next())
// is evaluated right-to-left by g++ (see oracle/vsnray_oracle_frame.c), so the first draw is u2.
VSB_HD V3 lambert_sample(const Material& m, V3 n, V3& wi, float& pdf, uint32_t& samp)
{
    V3 u, v;
    ortho_basis(u, v, n);
    const float u2 = sampler_next(samp);
    const float u1 = sampler_next(samp);
    wi = hemisphere_dir(u, v, n, cosine_sample_hemisphere(u1, u2));
    pdf = rn_mul(vdot(n, wi), VSB_INV_PI);
    return lambert_f(m);
}

// blinn::sample_f -- brdf.h:121-150
VSB_HD V3 blinn_sample(const Material& m, V3 n, V3 wo, V3& wi, float& pdf, uint32_t& samp)
{
    const float u1 = sampler_next(samp);
    const float u2 = sampler_next(samp);
    const float e = m.spec_exp;
    const float costheta = powf(u1, 1.0f / (e + 1.0f));
    const float sintheta = rn_sqrt(max0(rn_sub(1.0f, rn_mul(costheta, costheta))));
    const float phi = rn_mul(u2, VSB_TWO_PI);
    V3 u, v;
    ortho_basis(u, v, n);
    const V3 h = vnormalize(vadd(vadd(vscale(u, rn_mul(sintheta, cosf(phi))), vscale(v, rn_mul(sintheta, sinf(phi)))), vscale(n, costheta)));
    wi = vreflect(wo, h);
    const float vdoth = vdot(wo, h);
    pdf = ((e + 1.0f) * powf(costheta, e)) / (2.0f * VSB_PI * 4.0f * vdoth);
    return blinn_f(m, n, wo, wi);
}

// plastic::sample_impl -- detail/material/plastic.inl:186-229; matte -- matte.inl:147-155
VSB_HD V3 material_sample(const Material& m, V3 n, V3 wo, V3& wi, float& pdf, uint32_t& samp)
{
    if (m.kind == 1) return lambert_sample(m, n, wi, pdf, samp);
    float prob_diff = ((m.cd[0] + m.cd[1] + m.cd[2]) / 3.0f) * m.kd;
    float prob_spec = ((m.cs[0] + m.cs[1] + m.cs[2]) / 3.0f) * m.ks;
    if (prob_diff == 0.0f && prob_spec == 0.0f) { prob_diff = 0.5f; prob_spec = 0.5f; }
    prob_diff = prob_diff / (prob_diff + prob_spec);
    const float u = sampler_next(samp);
    if (u < prob_diff) return lambert_sample(m, n, wi, pdf, samp);
    return blinn_sample(m, n, wo, wi, pdf, samp);
}

// ---- the kernels; Tracer provides closest(o, d, hit) and occluded(o, d, max_t) ------------------

template <typename Tracer>
VSB_HD void kernel_simple(const FrameCtx& f, Tracer& tr, V3 o, V3 d, float out[4])
{
    HitReg h;
    tr.closest(o, d, h);
    if (!h.hit) { out[0] = f.p.bg[0]; out[1] = f.p.bg[1]; out[2] = f.p.bg[2]; out[3] = f.p.bg[3]; return; }
    const V3 pos = vadd(o, vscale(d, h.t));
    const Material m = f.materials[h.geom_id];
    const float4 nn = f.normals[h.prim_id];
    V3 n = v3(nn.x, nn.y, nn.z);
    const float aw = f.p.ambient[3];
    V3 shaded = v3(m.ca[0] * m.ka * (f.p.ambient[0] * aw), m.ca[1] * m.ka * (f.p.ambient[1] * aw), m.ca[2] * m.ka * (f.p.ambient[2] * aw));
    const V3 view = vneg(d);
    n = vfaceforward(n, view, n);
    for (int i = 0; i < f.p.n_lights; ++i)
    {
        const PointLight l = f.lights[i];
        const V3 wi = vnormalize(vsub(ld3(l.pos), pos));
        const V3 c = shade_light(m, l, pos, n, view, wi);
        shaded = v3(shaded.x + c.x, shaded.y + c.y, shaded.z + c.z);
    }
    out[0] = shaded.x; out[1] = shaded.y; out[2] = shaded.z; out[3] = 1.0f;
}

template <typename Tracer>
VSB_HD void kernel_whitted(const FrameCtx& f, Tracer& tr, V3 o, V3 d, float out[4])
{
    HitReg h;
    tr.closest(o, d, h);
    if (!h.hit) { out[0] = f.p.bg[0]; out[1] = f.p.bg[1]; out[2] = f.p.bg[2]; out[3] = f.p.bg[3]; return; }
    V3 color = v3(0.0f, 0.0f, 0.0f);
    unsigned depth = 0;
    float throughput = 1.0f;
    const float aw = f.p.ambient[3];
    while (h.hit && throughput > f.p.epsilon && depth++ < f.p.num_bounces)
    {
        const V3 pos = vadd(o, vscale(d, h.t));
        const Material m = f.materials[h.geom_id];
        const float4 nn = f.normals[h.prim_id];
        const V3 sn = v3(nn.x, nn.y, nn.z);
        V3 shaded = v3(m.ca[0] * m.ka * (f.p.ambient[0] * aw), m.ca[1] * m.ka * (f.p.ambient[1] * aw), m.ca[2] * m.ka * (f.p.ambient[2] * aw));
        const V3 view = vneg(d);
        const V3 n = vfaceforward(sn, view, sn);
        for (int i = 0; i < f.p.n_lights; ++i)
        {
            const PointLight l = f.lights[i];
            const V3 lpos = ld3(l.pos);
            const V3 wi = vnormalize(vsub(lpos, pos));
            const V3 so = vadd(pos, vscale(wi, f.p.epsilon));
            const float ld = vlength(vsub(pos, lpos));
            if (!tr.occluded(so, wi, ld))
            {
                const V3 c = shade_light(m, l, pos, n, view, wi);
                shaded = v3(shaded.x + c.x, shaded.y + c.y, shaded.z + c.z);
            }
        }
        color = v3(color.x + shaded.x * throughput, color.y + shaded.y * throughput, color.z + shaded.z * throughput);
        const float kr = (m.kind == 0) ? 0.1f : 0.0f;   // specular_bounce, detail/whitted.inl:79-110
        if (kr > 0.0f)
        {
            const V3 dir = vreflect(view, sn);
            o = vadd(pos, vscale(dir, f.p.epsilon));
            d = dir;
            tr.closest(o, d, h);
        }
        throughput = rn_mul(throughput, kr);
    }
    out[0] = color.x; out[1] = color.y; out[2] = color.z; out[3] = 1.0f;
}

template <typename Tracer>
VSB_HD void kernel_ao(const FrameCtx& f, Tracer& tr, V3 o, V3 d, uint32_t& samp, float out[4])
{
    HitReg h;
    tr.closest(o, d, h);
    if (!h.hit) { out[0] = f.p.bg[0]; out[1] = f.p.bg[1]; out[2] = f.p.bg[2]; out[3] = f.p.bg[3]; return; }
    const V3 pos = vadd(o, vscale(d, h.t));
    const float4 nn = f.normals[h.prim_id];
    const V3 w = v3(nn.x, nn.y, nn.z);
    V3 u, v;
    ortho_basis(u, v, w);
    float clr = 1.0f;
    const float dec = 1.0f / (float)f.p.ao_samples;
    for (int i = 0; i < f.p.ao_samples; ++i)
    {
        const float u2 = sampler_next(samp);   // right-to-left argument evaluation (see above)
        const float u1 = sampler_next(samp);
        const V3 dir = hemisphere_dir(u, v, w, cosine_sample_hemisphere(u1, u2));
        const V3 ao_o = vadd(pos, vscale(dir, 1E-3f));
        if (tr.occluded(ao_o, dir, f.p.ao_radius)) clr = rn_sub(clr, dec);
    }
    out[0] = clr; out[1] = clr; out[2] = clr; out[3] = 1.0f;
}

template <typename Tracer>
VSB_HD void kernel_pathtracing(const FrameCtx& f, Tracer& tr, V3 o, V3 d, uint32_t& samp, float out[4])
{
    bool active = true, result_hit = false;
    V3 dst = v3(1.0f, 1.0f, 1.0f);
    const float aw = f.p.ambient[3];
    for (unsigned bounce = 0; bounce < f.p.num_bounces; ++bounce)
    {
        HitReg h;
        tr.closest(o, d, h);
        if (active && !h.hit) dst = v3(dst.x * (f.p.ambient[0] * aw), dst.y * (f.p.ambient[1] * aw), dst.z * (f.p.ambient[2] * aw));
        active = active && h.hit;
        if (!active) break;
        if (bounce == 0) result_hit = true;
        const V3 view = vneg(d);
        const Material m = f.materials[h.geom_id];
        const float4 nn = f.normals[h.prim_id];
        const V3 sn = v3(nn.x, nn.y, nn.z);
        const V3 n = vfaceforward(sn, view, sn);
        V3 refl = v3(0.0f, 0.0f, 0.0f);
        float pdf = 0.0f;
        V3 src = material_sample(m, n, view, refl, pdf, samp);
        const bool zero_pdf = pdf <= 0.0f;
        const float w = vdot(n, refl) / pdf;
        if (!zero_pdf) dst = v3(dst.x * (src.x * w), dst.y * (src.y * w), dst.z * (src.z * w));
        else           dst = v3(0.0f, 0.0f, 0.0f);
        active = active && !zero_pdf;
        if (!active) break;
        const V3 pos = vadd(o, vscale(d, h.t));
        o = vadd(pos, vscale(refl, f.p.epsilon));
        d = refl;
    }
    if (active) dst = v3(0.0f, 0.0f, 0.0f);
    if (result_hit) { out[0] = dst.x; out[1] = dst.y; out[2] = dst.z; out[3] = 1.0f; }
    else { out[0] = f.p.bg[0]; out[1] = f.p.bg[1]; out[2] = f.p.bg[2]; out[3] = f.p.bg[3]; }
}

} // namespace vsb
