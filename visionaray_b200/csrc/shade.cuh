// shade.cuh -- device restatement of the frame() callers of the ray-query path: primary ray
// generation (detail/sched_common.h:130-150), the built-in kernels simple / whitted /
// pathtracing (detail/{simple,whitted,pathtracing}.inl) with the plastic / matte + point-light
// subset (detail/material/{plastic,matte}.inl, brdf.h, detail/point_light.inl), and the AO
// kernel of src/examples/ao/main.cpp:183-246.
//
// Everything that DEFINES A RAY (primary rays, hit positions, light / reflection / sampled
// directions, offsets) uses single-rounding arithmetic in the reference's operation order, so
// secondary rays are the same rays the oracle traces; sin / cos / pow of sampled directions are
// correctly rounded on both sides (ray_sincosf / ray_powf below), where the reference's libm is
// within one ulp.  Colour arithmetic is free to use FMAs: pixel parity is a 1/255 tolerance.
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

// unorm<8> <-> float (math/detail/unorm.inl:18-31): texels and 8-bit render targets
VSB_HD uint32_t float_to_unorm8(float f)        // math/detail/unorm.inl:18-24, after the clamp of color_conversion.h:224
{
    f = f < 0.0f ? 0.0f : (f > 1.0f ? 1.0f : f);
    return (uint32_t)((double)f * 255.0);
}
#if defined(__CUDACC__)
// float(u) / double(255) for u = 0..255 (math/detail/unorm.inl:26-31), tabulated by vsb_init(): the exact values
// without three or four inlined fp64 divisions per texel / pixel in programs that are sensitive to code size
__constant__ float c_unorm8_to_float[256];
#endif
VSB_HD float unorm8_to_float(uint32_t u)
{
#if defined(__CUDA_ARCH__)
    return c_unorm8_to_float[u & 255u];
#else
    return (float)((double)(float)u / 255.0);
#endif
}


struct Material { float ca[3], ka, cd[3], kd, cs[3], ks, spec_exp; int32_t kind; float pad[2]; };
struct PointLight { float pos[3], kl, cl[3], att_const, att_lin, att_quad; int32_t kind; float spot_exponent, spot_dir[3], spot_cos_cutoff; };

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

// texture_ref<vector<4, unorm<8>>, 2> on the device (texture/detail/texture2d.h): RGBA8, row-major
struct DevTexture { const uchar4* data; int32_t width, height, filter, addr_s, addr_t, pad; };

// What get_surface() reads besides the hit record (kernels.h:194-319, get_surface.h:330-530);
// attached to the scene by vsb_scene_set_surface().  enabled == 0: per-face normals only.
struct SurfCtx
{
    int32_t enabled, normal_binding, color_binding, n_textures;
    const float4*     normals;      // per face: [prim_id]; per vertex: [3 prim_id + k]
    const float4*     colors;       // per face / per vertex like the normals; null without colours
    const float2*     tex_coords;   // [3 prim_id + k]; null without textures
    const DevTexture* textures;     // by geom_id
    const float4*     ref_spheres;  // binding "none" on a sphere scene: the basic_sphere records (3 x 16 B) ...
    const uint32_t*   indices;      // ... and the index_bvh indirection (null for a plain bvh)
    const float4*     ref_generic;  // generic_primitive scene: the 80-byte variant records (5 x 16 B, `which` in word 16); a
                                    // sphere alternative takes get_normal(hr, sphere) under every binding
    const float4*     geo_normals;  // triangles, per vertex / no binding: normalize(cross(e1, e2)) of primitive(i), by primitive_list_index i
                                    // (precomputed by surface_geo_normals_kernel: one load at the hit instead of
                                    // indices[i] -> 64-byte triangle -> cross -> normalize)
};

struct FrameCtx
{
    FrameParams p;
    float eye[3], cu[3], cv[3], cw[3];
    const float4*     normals;     // per-face normals, 16 B each, by prim_id
    const Material*   materials;   // by geom_id
    const PointLight* lights;
    // sched_params built from matrices (camera_kind == 1): column-major matrix<4,4,float>
    int   camera_kind;
    float view[16], inv_view[16], proj[16], inv_proj[16];
    SurfCtx surf;
};

VSB_HD V3 ld3(const float* p) { return v3(p[0], p[1], p[2]); }

// matrix<4,4> * vector<4>, row r -- math/detail/matrix4.inl:171-181
VSB_HD float mat4_row(const float* m, int r, float x, float y, float z, float w)
{
    return rn_add(rn_add(rn_add(rn_mul(m[r], x), rn_mul(m[4 + r], y)), rn_mul(m[8 + r], z)), rn_mul(m[12 + r], w));
}
VSB_HD void mat4_mul(const float* m, const float v[4], float out[4])
{
    for (int r = 0; r < 4; ++r) out[r] = mat4_row(m, r, v[0], v[1], v[2], v[3]);
}

// make_primary_ray_impl, matrix overload: detail/sched_common.h:152-178
VSB_HD void primary_ray_matrix(const FrameCtx& f, float x, float y, V3& ori, V3& dir)
{
    const float u = rn_sub(rn_div(rn_mul(2.0f, rn_add(x, 0.5f)), (float)f.p.width), 1.0f);
    const float v = rn_sub(rn_div(rn_mul(2.0f, rn_add(y, 0.5f)), (float)f.p.height), 1.0f);
    const float n4[4] = { u, v, -1.0f, 1.0f }, f4[4] = { u, v, 1.0f, 1.0f };
    float t[4], o[4], d[4];
    mat4_mul(f.inv_proj, n4, t); mat4_mul(f.inv_view, t, o);
    mat4_mul(f.inv_proj, f4, t); mat4_mul(f.inv_view, t, d);
    ori = v3(rn_div(o[0], o[3]), rn_div(o[1], o[3]), rn_div(o[2], o[3]));
    dir = vnormalize(vsub(v3(rn_div(d[0], d[3]), rn_div(d[1], d[3]), rn_div(d[2], d[3])), ori));
}

// depth_transform: detail/sched_common.h:337-354
VSB_HD float depth_transform(const FrameCtx& f, V3 pos)
{
    const float p4[4] = { pos.x, pos.y, pos.z, 1.0f };
    float t[4], q[4];
    mat4_mul(f.view, p4, t); mat4_mul(f.proj, t, q);
    return rn_mul(rn_add(rn_div(q[2], q[3]), 1.0f), 0.5f);
}

// make_primary_ray_impl, pinhole: detail/sched_common.h:130-150
VSB_HD void primary_ray(const FrameCtx& f, float x, float y, V3& ori, V3& dir)
{
    const float u = rn_sub(rn_div(rn_mul(2.0f, rn_add(x, 0.5f)), (float)f.p.width), 1.0f);
    const float v = rn_sub(rn_div(rn_mul(2.0f, rn_add(y, 0.5f)), (float)f.p.height), 1.0f);
    ori = ld3(f.eye);
    dir = vnormalize(vadd(vadd(vscale(ld3(f.cu), u), vscale(ld3(f.cv), v)), ld3(f.cw)));
}

// point_light::intensity -- detail/point_light.inl:11-27
// SPOT: the instantiation handles spot lights (the lean frame programs are point-light only: the cone costs the
// cfg2 frame 2 % even when no light uses it)
template <bool SPOT>
VSB_HD V3 light_intensity(const PointLight& l, V3 pos)
{
    const float dist = vlength(vsub(ld3(l.pos), pos));
    float att = 1.0f / (l.att_const + l.att_lin * dist + l.att_quad * dist * dist);
    if (SPOT && l.kind == 1)
    {
        // spot_light::intensity -- detail/spot_light.inl:14-38
        const V3 to = vnormalize(vsub(pos, ld3(l.pos)));
        const float spot = vdot(ld3(l.spot_dir), to);
        att *= spot > l.spot_cos_cutoff ? powf(spot, l.spot_exponent) : 0.0f;
    }
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
    // pow(1 - sat, 5) of brdf.h:113 as four multiplications: within 2 ulp of powf, and ~150 instructions less in
    // every frame program (colour arithmetic: pixel parity is a 1/255 tolerance)
    const float om = 1.0f - sat, om2 = om * om;
    const float sch = om2 * om2 * om;
    const float nfactor = (m.spec_exp + 2.0f) / (8.0f * VSB_PI);
    // pow(hdotn, exp) with hdotn in [0, 1]: exp2(exp * log2(hdotn)) on the SFU (relative error ~1e-6 * exp, far
    // inside the 1/255 pixel tolerance) instead of the ~150-instruction powf; 0^0 = 1 kept
#if defined(__CUDA_ARCH__)
    const float ph = (m.spec_exp == 0.0f ? 1.0f : __powf(hdotn, m.spec_exp)) * nfactor;
#else
    const float ph = powf(hdotn, m.spec_exp) * nfactor;
#endif
    const float sx = m.cs[0] * m.ks, sy = m.cs[1] * m.ks, sz = m.cs[2] * m.ks;
    return v3((sx + (1.0f - sx) * sch) * ph, (sy + (1.0f - sy) * sch) * ph, (sz + (1.0f - sz) * sch) * ph);
}

// plastic::shade / matte::shade -- detail/material/plastic.inl:18-37, matte.inl:18-39
VSB_HD V3 shade_light(const Material& m, const PointLight& l, V3 pos, V3 n, V3 wo, V3 wi)
{
    const float ndotl = max0(vdot(n, wi));
    V3 f = lambert_f(m);
    if (m.kind == 0) { const V3 s = blinn_f(m, n, wo, wi); f = v3(f.x + s.x, f.y + s.y, f.z + s.z); }
    const V3 li = light_intensity<false>(l, pos);
    return v3(f.x * VSB_PI * li.x * ndotl, f.y * VSB_PI * li.y * ndotl, f.z * VSB_PI * li.z * ndotl);
}

// make_orthonormal_basis -- math/detail/vector3.inl:357-367
VSB_HD void ortho_basis(V3& u, V3& v, V3 w)
{
    if (fabsf(w.x) > fabsf(w.y)) v = vnormalize(v3(-w.z, 0.0f, w.x));
    else                         v = vnormalize(v3(0.0f, w.z, -w.y));
    u = vcross(v, w);
}

// ---- sin / cos / pow of the values that define a sampled direction --------------------------------
// The reference calls std::sin / std::cos / std::pow (libm).  CUDA's sinf / cosf / powf are a few ulps
// away from libm's, which forks a path at a triangle edge once in a few thousand bounces; so these three
// are evaluated in double with a FIXED sequence of single-rounding operations and rounded to float once
// -- the same sequence as the oracle's (oracle/vso_raymath.h restates it in C), i.e. the same bits, and
// the correctly rounded value in all but ~2^-22 of arguments.
#if defined(__CUDA_ARCH__)
VSB_D double d_mul(double a, double b) { return __dmul_rn(a, b); }
VSB_D double d_add(double a, double b) { return __dadd_rn(a, b); }
VSB_D double d_sub(double a, double b) { return __dsub_rn(a, b); }
VSB_D double d_div(double a, double b) { return __ddiv_rn(a, b); }
VSB_D double bits_to_double(uint64_t b) { return __longlong_as_double((long long)b); }
VSB_D uint64_t double_to_bits(double d) { return (uint64_t)__double_as_longlong(d); }
#else
VSB_HD double d_mul(double a, double b) { return a * b; }            // host build: -ffp-contract=off
VSB_HD double d_add(double a, double b) { return a + b; }
VSB_HD double d_sub(double a, double b) { return a - b; }
VSB_HD double d_div(double a, double b) { return a / b; }
VSB_HD double bits_to_double(uint64_t b) { double d; memcpy(&d, &b, 8); return d; }
VSB_HD uint64_t double_to_bits(double d) { uint64_t b; memcpy(&b, &d, 8); return b; }
#endif
VSB_HD double d_horner(double p, double x, double c) { return d_add(d_mul(p, x), c); }

// |x| < 2^20
VSB_HD void ray_sincosf(float x, float& s, float& c)
{
    const double PIO2_HI = 1.5707963267341256e+00, PIO2_LO = 6.0771005065061922e-11, TWO_OVER_PI = 6.3661977236758138e-01;
    const double xd = (double)x;
    const double kd = rint(d_mul(xd, TWO_OVER_PI));
    const double r = d_sub(d_sub(xd, d_mul(kd, PIO2_HI)), d_mul(kd, PIO2_LO));
    const double r2 = d_mul(r, r);
    double ps = -1.0 / 355687428096000.0;
    ps = d_horner(ps, r2, 1.0 / 1307674368000.0);
    ps = d_horner(ps, r2, -1.0 / 6227020800.0);
    ps = d_horner(ps, r2, 1.0 / 39916800.0);
    ps = d_horner(ps, r2, -1.0 / 362880.0);
    ps = d_horner(ps, r2, 1.0 / 5040.0);
    ps = d_horner(ps, r2, -1.0 / 120.0);
    ps = d_horner(ps, r2, 1.0 / 6.0);
    const double sn = d_sub(r, d_mul(d_mul(r, r2), ps));
    double pc = 1.0 / 6402373705728000.0;
    pc = d_horner(pc, r2, -1.0 / 20922789888000.0);
    pc = d_horner(pc, r2, 1.0 / 87178291200.0);
    pc = d_horner(pc, r2, -1.0 / 479001600.0);
    pc = d_horner(pc, r2, 1.0 / 3628800.0);
    pc = d_horner(pc, r2, -1.0 / 40320.0);
    pc = d_horner(pc, r2, 1.0 / 720.0);
    pc = d_horner(pc, r2, -1.0 / 24.0);
    pc = d_horner(pc, r2, 0.5);
    const double cs = d_sub(1.0, d_mul(r2, pc));
    const int q = (int)((long long)kd & 3);
    const double so = (q & 1) ? cs : sn, co = (q & 1) ? sn : cs;
    s = (float)((q & 2) ? -so : so);
    c = (float)(((q + 1) & 2) ? -co : co);
}

// x^y, x >= 0 (x < 0 -> NaN)
VSB_HD float ray_powf(float x, float y)
{
    const double LN2_HI = 6.9314718036912382e-01, LN2_LO = 1.9082149292705877e-10, INV_LN2 = 1.4426950408889634e+00;
    if (y == 0.0f) return 1.0f;
    if (x != x || y != y || x < 0.0f) return u2f(0x7fc00000u);
    if (x == 0.0f) return y > 0.0f ? 0.0f : u2f(0x7f800000u);
    if (x == 1.0f) return 1.0f;
    if (x == u2f(0x7f800000u)) return y > 0.0f ? x : 0.0f;
    uint64_t b = double_to_bits((double)x);
    int k = (int)((b >> 52) & 0x7ffu) - 1023;
    double m = bits_to_double((b & 0x000fffffffffffffull) | 0x3ff0000000000000ull);
    if (m > 1.4142135623730951) { m = d_mul(m, 0.5); k = k + 1; }
    const double s = d_div(d_sub(m, 1.0), d_add(m, 1.0));
    const double s2 = d_mul(s, s);
    double p = 1.0 / 21.0;
    p = d_horner(p, s2, 1.0 / 19.0);
    p = d_horner(p, s2, 1.0 / 17.0);
    p = d_horner(p, s2, 1.0 / 15.0);
    p = d_horner(p, s2, 1.0 / 13.0);
    p = d_horner(p, s2, 1.0 / 11.0);
    p = d_horner(p, s2, 1.0 / 9.0);
    p = d_horner(p, s2, 1.0 / 7.0);
    p = d_horner(p, s2, 1.0 / 5.0);
    p = d_horner(p, s2, 1.0 / 3.0);
    const double ts = d_mul(2.0, s);
    const double lnm = d_add(ts, d_mul(ts, d_mul(s2, p)));
    const double kd = (double)k;
    const double lnx = d_add(d_mul(kd, LN2_HI), d_add(d_mul(kd, LN2_LO), lnm));
    const double z = d_mul((double)y, lnx);
    if (z > 89.0) return u2f(0x7f800000u);
    if (z < -104.0) return 0.0f;
    const double nd = rint(d_mul(z, INV_LN2));
    const double r = d_sub(d_sub(z, d_mul(nd, LN2_HI)), d_mul(nd, LN2_LO));
    double e = 1.0 / 6227020800.0;
    e = d_horner(e, r, 1.0 / 479001600.0);
    e = d_horner(e, r, 1.0 / 39916800.0);
    e = d_horner(e, r, 1.0 / 3628800.0);
    e = d_horner(e, r, 1.0 / 362880.0);
    e = d_horner(e, r, 1.0 / 40320.0);
    e = d_horner(e, r, 1.0 / 5040.0);
    e = d_horner(e, r, 1.0 / 720.0);
    e = d_horner(e, r, 1.0 / 120.0);
    e = d_horner(e, r, 1.0 / 24.0);
    e = d_horner(e, r, 1.0 / 6.0);
    e = d_horner(e, r, 0.5);
    e = d_horner(e, r, 1.0);
    e = d_horner(e, r, 1.0);
    const double sc = bits_to_double((uint64_t)((long long)nd + 1023) << 52);
    return (float)d_mul(e, sc);
}

// cosine_sample_hemisphere -- sampling.h:61-71
VSB_HD V3 cosine_sample_hemisphere(float u1, float u2)
{
    const float r = rn_sqrt(u1);
    const float theta = rn_mul(VSB_TWO_PI, u2);
    float sn, cs;
    ray_sincosf(theta, sn, cs);
    return v3(rn_mul(r, cs), rn_mul(r, sn), rn_sqrt(max0(rn_sub(1.0f, u1))));
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
    const float costheta = ray_powf(u1, rn_div(1.0f, rn_add(e, 1.0f)));
    const float sintheta = rn_sqrt(max0(rn_sub(1.0f, rn_mul(costheta, costheta))));
    const float phi = rn_mul(u2, VSB_TWO_PI);
    V3 u, v;
    ortho_basis(u, v, n);
    float sp, cp;
    ray_sincosf(phi, sp, cp);
    const V3 h = vnormalize(vadd(vadd(vscale(u, rn_mul(sintheta, cp)), vscale(v, rn_mul(sintheta, sp))), vscale(n, costheta)));
    wi = vreflect(wo, h);
    const float vdoth = vdot(wo, h);
    pdf = ((e + 1.0f) * powf(costheta, e)) / (2.0f * VSB_PI * 4.0f * vdoth);
    return blinn_f(m, n, wo, wi);
}

// plastic::sample_impl -- detail/material/plastic.inl:186-229; matte -- matte.inl:147-155
VSB_HD V3 material_sample(const Material& m, V3 n, V3 wo, V3& wi, float& pdf, uint32_t& samp)
{
    // matte always takes the diffuse lobe; plastic draws u and picks (one copy of each lobe's code)
    bool diffuse = true;
    if (m.kind != 1)
    {
        float prob_diff = ((m.cd[0] + m.cd[1] + m.cd[2]) / 3.0f) * m.kd;
        float prob_spec = ((m.cs[0] + m.cs[1] + m.cs[2]) / 3.0f) * m.ks;
        if (prob_diff == 0.0f && prob_spec == 0.0f) { prob_diff = 0.5f; prob_spec = 0.5f; }
        prob_diff = prob_diff / (prob_diff + prob_spec);
        diffuse = sampler_next(samp) < prob_diff;
    }
    if (diffuse) return lambert_sample(m, n, wi, pdf, samp);
    return blinn_sample(m, n, wo, wi, pdf, samp);
}

// ---- get_surface() with bindings, colours and textures; generic_material -------------------------
//
// Used by the SURF instantiations of the frame programs (kernel_params that carry more than
// per-face normals and a plastic / matte list).  Normals and everything else that defines a ray
// keep the reference's operation order with single roundings; so does the texture filter, because
// its result is TRUNCATED to an integer texel value before normalisation (texture/detail/
// sampler2d.h:81-113), which would turn an ulp into 1/255.

struct Surface { V3 gn, sn, tex; bool has_tex; };

// lerp(a, b, c, u, v) = (a (1 - (u + v)) + c v) + b u -- math/detail/math.h:466-475
VSB_HD V3 vlerp3(V3 a, V3 b, V3 c, float u, float v)
{
    return vadd(vadd(vscale(a, rn_sub(1.0f, rn_add(u, v))), vscale(c, v)), vscale(b, u));
}
VSB_HD float lerp1(float a, float b, float x) { return rn_add(rn_mul(rn_sub(1.0f, x), a), rn_mul(x, b)); }   // :459-464

// map_tex_coord -- texture/detail/filter/common.h:24-48 (0 Wrap, 1 Mirror, 2 Clamp)
VSB_HD float map_tex_coord(float coord, int texsize, int mode)
{
    if (mode == 0) return rn_sub(coord, floorf(coord));
    if (mode == 1)
    {
        const float fl = floorf(coord);
        if ((((int)fl) & 1) == 1) return rn_sub(rn_div((float)(texsize - 1), (float)texsize), rn_sub(coord, fl));
        return rn_sub(coord, fl);
    }
    const float hi = rn_sub(1.0f, rn_div(1.0f, (float)texsize));
    return coord < 0.0f ? 0.0f : (coord > hi ? hi : coord);
}

#if defined(__CUDACC__)
__device__ __forceinline__ uchar4 texel(const DevTexture& t, int x, int y)
{
    // the reference reads out of bounds when a wrapped coordinate rounds up to 1 (coord - floor(coord)
    // == 1 for tiny negative coord); the index is clamped here instead
    x = x < 0 ? 0 : (x >= t.width ? t.width - 1 : x);
    y = y < 0 ? 0 : (y >= t.height ? t.height - 1 : y);
    return __ldg(t.data + (size_t)y * (size_t)t.width + (size_t)x);
}

// tex2D(texture_ref<vector<4, unorm<8>>, 2>, coord).xyz -- texture/detail/sampler2d.h:81-151,
// filter/nearest.h:45-66, filter/linear.h:61-98, math/detail/unorm.inl:26-31
// ---- BSpline / CardinalSpline filters (texture/detail/filter/{cubic_opt,cubic}.h, weights filter/common.h:104-192).
// Rare and bulky, so ONE out-of-line function that tex2d_rgb() only calls for those filter modes.  The weights mix
// float products with double literals exactly as the reference's expressions do (single roundings spelled out: the
// filtered value is truncated to an integer texel value afterwards).
__device__ __forceinline__ double dmul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double dadd(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float a3f(float a) { return rn_mul(rn_mul(a, a), a); }                      // a * a * a in float
__device__ __forceinline__ double a3d(double c, float a) { return dmul(dmul(dmul(c, a), a), a); }       // c * a * a * a in double
__device__ __forceinline__ double a2d(double c, float a) { return dmul(dmul(c, a), a); }                // c * a * a
__device__ __forceinline__ float bs_w0(float a) { return (float)dmul(1.0 / 6.0, dadd(dadd(dadd((double)(-a3f(a)), a2d(3.0, a)), -dmul(3.0, a)), 1.0)); }
__device__ __forceinline__ float bs_w1(float a) { return (float)dmul(1.0 / 6.0, dadd(dadd(a3d(3.0, a), -a2d(6.0, a)), 4.0)); }
__device__ __forceinline__ float bs_w2(float a) { return (float)dmul(1.0 / 6.0, dadd(dadd(dadd(a3d(-3.0, a), a2d(3.0, a)), dmul(3.0, a)), 1.0)); }
__device__ __forceinline__ float bs_w3(float a) { return (float)dmul(1.0 / 6.0, (double)a3f(a)); }
__device__ __forceinline__ float cs_w0(float a) { return (float)dadd(dadd(a3d(-0.5, a), (double)rn_mul(a, a)), -dmul(0.5, a)); }
__device__ __forceinline__ float cs_w1(float a) { return (float)dadd(dadd(a3d(1.5, a), -a2d(2.5, a)), 1.0); }
__device__ __forceinline__ float cs_w2(float a) { return (float)dadd(dadd(a3d(-1.5, a), a2d(2.0, a)), dmul(0.5, a)); }
__device__ __forceinline__ float cs_w3(float a) { return (float)dadd(a3d(0.5, a), -a2d(0.5, a)); }

__device__ __forceinline__ void tex_linear_q(const DevTexture& t, float s, float tt, int q[3])
{
    const float fw = (float)t.width, fh = (float)t.height;
    const float hs = rn_div(0.5f, fw), ht = rn_div(0.5f, fh);
    const float s1 = map_tex_coord(rn_sub(s, hs), t.width, t.addr_s), t1 = map_tex_coord(rn_sub(tt, ht), t.height, t.addr_t);
    const float s2 = map_tex_coord(rn_add(s, hs), t.width, t.addr_s), t2 = map_tex_coord(rn_add(tt, ht), t.height, t.addr_t);
    const int lx = __float2int_rz(rn_mul(s1, fw)), ly = __float2int_rz(rn_mul(t1, fh));
    const int hx = __float2int_rz(rn_mul(s2, fw)), hy = __float2int_rz(rn_mul(t2, fh));
    const uchar4 e00 = texel(t, lx, ly), e10 = texel(t, hx, ly), e01 = texel(t, lx, hy), e11 = texel(t, hx, hy);
    const float u = rn_sub(rn_mul(s1, fw), (float)lx), v = rn_sub(rn_mul(t1, fh), (float)ly);
    q[0] = __float2int_rz(lerp1(lerp1((float)e00.x, (float)e10.x, u), lerp1((float)e01.x, (float)e11.x, u), v));
    q[1] = __float2int_rz(lerp1(lerp1((float)e00.y, (float)e10.y, u), lerp1((float)e01.y, (float)e11.y, u), v));
    q[2] = __float2int_rz(lerp1(lerp1((float)e00.z, (float)e10.z, u), lerp1((float)e01.z, (float)e11.z, u), v));
}

__device__ __noinline__ V3 tex2d_cubic_rgb(const DevTexture& t, float s, float tt)
{
    const float fw = (float)t.width, fh = (float)t.height;
    int q[3];
    if (t.filter == 2)
    {
        // cubic_opt: four bilinear lookups at the g / h offsets (filter/cubic_opt.h:58-95)
        const float x = rn_sub(rn_mul(s, fw), 0.5f), floorx = floorf(x), fracx = rn_sub(x, floorx);
        const float y = rn_sub(rn_mul(tt, fh), 0.5f), floory = floorf(y), fracy = rn_sub(y, floory);
        const float w0x = bs_w0(fracx), w1x = bs_w1(fracx), w2x = bs_w2(fracx), w3x = bs_w3(fracx);
        const float w0y = bs_w0(fracy), w1y = bs_w1(fracy), w2y = bs_w2(fracy), w3y = bs_w3(fracy);
        const float g0x = rn_add(w0x, w1x), g1x = rn_add(w2x, w3x), g0y = rn_add(w0y, w1y), g1y = rn_add(w2y, w3y);
        const float h_00 = rn_div(rn_add(rn_sub(floorx, 0.5f), rn_div(w1x, g0x)), fw);
        const float h_10 = rn_div(rn_add(rn_add(floorx, 1.5f), rn_div(w3x, g1x)), fw);
        const float h_01 = rn_div(rn_add(rn_sub(floory, 0.5f), rn_div(w1y, g0y)), fh);
        const float h_11 = rn_div(rn_add(rn_add(floory, 1.5f), rn_div(w3y, g1y)), fh);
        int a[3], b[3], c[3], d[3];
        tex_linear_q(t, h_00, h_01, a); tex_linear_q(t, h_10, h_01, b); tex_linear_q(t, h_00, h_11, c); tex_linear_q(t, h_10, h_11, d);
        for (int k = 0; k < 3; ++k)
        {
            const float f_0 = rn_add(rn_mul(g0x, (float)a[k]), rn_mul(g1x, (float)b[k]));
            const float f_1 = rn_add(rn_mul(g0x, (float)c[k]), rn_mul(g1x, (float)d[k]));
            q[k] = __float2int_rz(rn_add(rn_mul(g0y, f_0), rn_mul(g1y, f_1)));
        }
    }
    else
    {
        // cubic with the cardinal-spline weights: 4 x 4 texels (filter/cubic.h:85-140)
        const float off[4] = { 1.5f, 0.5f, 0.5f, 1.5f };
        float cx[4], cy[4]; int px[4], py[4];
        for (int i = 0; i < 4; ++i)
        {
            const float ds = rn_div(off[i], fw), dt = rn_div(off[i], fh);
            cx[i] = map_tex_coord(i < 2 ? rn_sub(s, ds) : rn_add(s, ds), t.width, t.addr_s);
            cy[i] = map_tex_coord(i < 2 ? rn_sub(tt, dt) : rn_add(tt, dt), t.height, t.addr_t);
            px[i] = __float2int_rz(rn_mul(cx[i], fw)); py[i] = __float2int_rz(rn_mul(cy[i], fh));
        }
        const float ux = rn_sub(rn_mul(cx[1], fw), (float)px[1]), uy = rn_sub(rn_mul(cy[1], fh), (float)py[1]);
        const float wx[4] = { cs_w0(ux), cs_w1(ux), cs_w2(ux), cs_w3(ux) }, wy[4] = { cs_w0(uy), cs_w1(uy), cs_w2(uy), cs_w3(uy) };
        float acc[3] = { 0.0f, 0.0f, 0.0f };
        for (int j = 0; j < 4; ++j)
        {
            const uchar4 e0 = texel(t, px[0], py[j]), e1 = texel(t, px[1], py[j]), e2 = texel(t, px[2], py[j]), e3 = texel(t, px[3], py[j]);
            const float fj[3] = {
                rn_add(rn_add(rn_add(rn_mul(wx[0], (float)e0.x), rn_mul(wx[1], (float)e1.x)), rn_mul(wx[2], (float)e2.x)), rn_mul(wx[3], (float)e3.x)),
                rn_add(rn_add(rn_add(rn_mul(wx[0], (float)e0.y), rn_mul(wx[1], (float)e1.y)), rn_mul(wx[2], (float)e2.y)), rn_mul(wx[3], (float)e3.y)),
                rn_add(rn_add(rn_add(rn_mul(wx[0], (float)e0.z), rn_mul(wx[1], (float)e1.z)), rn_mul(wx[2], (float)e2.z)), rn_mul(wx[3], (float)e3.z)) };
            // w0 f0 + w1 f1 + w2 f2 + w3 f3, left to right
            for (int k = 0; k < 3; ++k) acc[k] = j == 0 ? rn_mul(wy[0], fj[k]) : rn_add(acc[k], rn_mul(wy[j], fj[k]));
        }
        for (int k = 0; k < 3; ++k) q[k] = __float2int_rz(acc[k]);
    }
    // a spline can leave [0, 255]: normalised through uint32 like the reference, not through the 8-bit table
    return v3((float)((double)(float)(uint32_t)q[0] / 255.0), (float)((double)(float)(uint32_t)q[1] / 255.0),
              (float)((double)(float)(uint32_t)q[2] / 255.0));
}

__device__ inline V3 tex2d_rgb(const DevTexture& t, float s, float tt)
{
    if (t.filter >= 2) return tex2d_cubic_rgb(t, s, tt);
    const float fw = (float)t.width, fh = (float)t.height;
    int q[3];
    if (t.filter == 0)
    {
        const float cs = map_tex_coord(s, t.width, t.addr_s), ct = map_tex_coord(tt, t.height, t.addr_t);
        const uchar4 e = texel(t, __float2int_rz(rn_mul(cs, fw)), __float2int_rz(rn_mul(ct, fh)));
        q[0] = e.x; q[1] = e.y; q[2] = e.z;
    }
    else
    {
        const float hs = rn_div(0.5f, fw), ht = rn_div(0.5f, fh);
        const float s1 = map_tex_coord(rn_sub(s, hs), t.width, t.addr_s), t1 = map_tex_coord(rn_sub(tt, ht), t.height, t.addr_t);
        const float s2 = map_tex_coord(rn_add(s, hs), t.width, t.addr_s), t2 = map_tex_coord(rn_add(tt, ht), t.height, t.addr_t);
        const int lx = __float2int_rz(rn_mul(s1, fw)), ly = __float2int_rz(rn_mul(t1, fh));
        const int hx = __float2int_rz(rn_mul(s2, fw)), hy = __float2int_rz(rn_mul(t2, fh));
        const uchar4 e00 = texel(t, lx, ly), e10 = texel(t, hx, ly), e01 = texel(t, lx, hy), e11 = texel(t, hx, hy);
        const float u = rn_sub(rn_mul(s1, fw), (float)lx), v = rn_sub(rn_mul(t1, fh), (float)ly);
        q[0] = __float2int_rz(lerp1(lerp1((float)e00.x, (float)e10.x, u), lerp1((float)e01.x, (float)e11.x, u), v));
        q[1] = __float2int_rz(lerp1(lerp1((float)e00.y, (float)e10.y, u), lerp1((float)e01.y, (float)e11.y, u), v));
        q[2] = __float2int_rz(lerp1(lerp1((float)e00.z, (float)e10.z, u), lerp1((float)e01.z, (float)e11.z, u), v));
    }
    return v3(unorm8_to_float((uint32_t)q[0]), unorm8_to_float((uint32_t)q[1]), unorm8_to_float((uint32_t)q[2]));
}

// get_normal(hr, triangle) = normalize(cross(e1, e2)) (get_normal.h:96-101) of primitive(i) =
// prims[indices[i]] (bvh.h:226-229) for every position i of the primitive list, same single-rounding
// arithmetic as at shading time
__global__ void surface_geo_normals_kernel(int prim_kind, const void* __restrict__ prims, const uint32_t* __restrict__ indices,
                                           size_t n_list, float4* __restrict__ out)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_list) return;
    const size_t pi = indices ? indices[i] : i;
    const float4* rec = reinterpret_cast<const float4*>(reinterpret_cast<const char*>(prims) + pi * prim_stride(prim_kind));
    if (!prim_is_triangle(prim_kind, rec)) { out[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f); return; }   // a sphere of a generic scene
    const float4 e1 = rec[2], e2 = rec[3];
    const V3 g = vnormalize(vcross(v3(e1.x, e1.y, e1.z), v3(e2.x, e2.y, e2.z)));
    out[i] = make_float4(g.x, g.y, g.z, 0.0f);
}

// get_surface(hit_rec, params) for an index_bvh of triangles -- get_surface.h:330-530
// (inlined at its single call site per program; as an out-of-line call -- measured -- the path tracer, which
// evaluates it at every bounce, lost 18 %: the 10-word result goes through local memory)
__device__ __forceinline__ Surface surface_at(const FrameCtx& f, uint32_t prim_id, uint32_t geom_id, uint32_t list_index, float u, float v, V3 pos)
{
    const SurfCtx& sf = f.surf;
    Surface S;
    S.has_tex = false; S.tex = v3(1.0f, 1.0f, 1.0f);
    if (!sf.enabled)
    {
        const float4 nn = __ldg(f.normals + prim_id);
        S.gn = S.sn = v3(nn.x, nn.y, nn.z);
        return S;
    }
    bool sphere_alternative = false;
    if (sf.ref_generic)
    {
        // detail/generic_primitive_get_normal.h: the visitor applies the rule of the alternative the variant holds
        const size_t pi = sf.indices ? __ldg(sf.indices + list_index) : list_index;
        const float4* rec = sf.ref_generic + pi * 5;
        if (__float_as_uint(__ldg(rec + 4).x) != 1u)
        {
            const float4 c = __ldg(rec + 1), r = __ldg(rec + 2);
            S.gn = S.sn = v3(rn_div(rn_sub(pos.x, c.x), r.x), rn_div(rn_sub(pos.y, c.y), r.x), rn_div(rn_sub(pos.z, c.z), r.x));
            sphere_alternative = true;
        }
    }
    if (sphere_alternative) { }
    else if (sf.normal_binding == 0)
    {
        const float4 nn = __ldg(sf.normals + prim_id);
        S.gn = S.sn = v3(nn.x, nn.y, nn.z);
    }
    else if (sf.normal_binding == 2)
    {
        // kernel_params without normals: get_normal(hr, primitive) (get_surface.h:133-150, get_normal.h:96-118)
        if (sf.ref_spheres)
        {
            const size_t pi = sf.indices ? __ldg(sf.indices + list_index) : list_index;
            const float4 c = __ldg(sf.ref_spheres + pi * 3 + 1), r = __ldg(sf.ref_spheres + pi * 3 + 2);
            S.gn = S.sn = v3(rn_div(rn_sub(pos.x, c.x), r.x), rn_div(rn_sub(pos.y, c.y), r.x), rn_div(rn_sub(pos.z, c.z), r.x));
        }
        else
        {
            const float4 g = __ldg(sf.geo_normals + list_index);
            S.gn = S.sn = v3(g.x, g.y, g.z);
        }
    }
    else
    {
        // geometric normal of primitive(primitive_list_index) (get_surface.h:313-327, get_normal.h:96-101),
        // shading normal interpolated (get_shading_normal.h:60-76)
        const float4 g = __ldg(sf.geo_normals + list_index);
        S.gn = v3(g.x, g.y, g.z);
        const float4* n = sf.normals + (size_t)prim_id * 3;
        const float4 a = __ldg(n), b = __ldg(n + 1), c = __ldg(n + 2);
        S.sn = vnormalize(vlerp3(v3(a.x, a.y, a.z), v3(b.x, b.y, b.z), v3(c.x, c.y, c.z), u, v));
    }
    if (sf.color_binding || sf.tex_coords)
    {
        S.has_tex = true;
        V3 tc = v3(1.0f, 1.0f, 1.0f);
        if (sf.tex_coords && (int)geom_id < sf.n_textures)
        {
            const DevTexture t = sf.textures[geom_id];
            if (t.width > 0 && t.height > 0)
            {
                const float2* q = sf.tex_coords + (size_t)prim_id * 3;
                const float2 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
                const float w0 = rn_sub(1.0f, rn_add(u, v));
                tc = tex2d_rgb(t, rn_add(rn_add(rn_mul(a.x, w0), rn_mul(c.x, v)), rn_mul(b.x, u)),
                                  rn_add(rn_add(rn_mul(a.y, w0), rn_mul(c.y, v)), rn_mul(b.y, u)));
            }
        }
        if (sf.color_binding == 1)
        {
            const float4 c = __ldg(sf.colors + prim_id);
            tc = v3(c.x * tc.x, c.y * tc.y, c.z * tc.z);
        }
        else if (sf.color_binding == 2)
        {
            const float4* q = sf.colors + (size_t)prim_id * 3;
            const float4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
            const V3 col = vlerp3(v3(a.x, a.y, a.z), v3(b.x, b.y, b.z), v3(c.x, c.y, c.z), u, v);
            tc = v3(col.x * tc.x, col.y * tc.y, col.z * tc.z);
        }
        S.tex = tc;
    }
    return S;
}
#endif // __CUDACC__

// generic_material<emissive, matte, mirror, plastic>::shade -- detail/generic_material.inl:36-43 ->
// plastic.inl:18-37,176-183; matte.inl:18-39,136-142; emissive.inl:22-27,76-90; mirror.inl:22-27
VSB_HD V3 shade_light_ext(const Material& m, const Surface& S, const PointLight& l, V3 pos, V3 n, V3 wo, V3 wi)
{
    if (m.kind == 2) return v3(S.tex.x * m.cd[0] * m.kd, S.tex.y * m.cd[1] * m.kd, S.tex.z * m.cd[2] * m.kd);   // tex == 1 without colours
    if (m.kind == 3) return v3(0.0f, 0.0f, 0.0f);
    const float ndotl = max0(vdot(n, wi));
    V3 f = lambert_f(m);
    f = v3(f.x * S.tex.x, f.y * S.tex.y, f.z * S.tex.z);
    if (m.kind == 0) { const V3 s = blinn_f(m, n, wo, wi); f = v3(f.x + s.x, f.y + s.y, f.z + s.z); }
    const V3 li = light_intensity<true>(l, pos);
    return v3(f.x * VSB_PI * li.x * ndotl, f.y * VSB_PI * li.y * ndotl, f.z * VSB_PI * li.z * ndotl);
}

// specular_bounce() -- detail/whitted.inl:66-122: matte 0, mirror kr, everything else 0.1
VSB_HD float specular_kr(const Material& m) { return m.kind == 1 ? 0.0f : (m.kind == 3 ? m.ks : 0.1f); }

// fresnel_reflectance(conductor_tag, eta, k, cosi) -- fresnel.h:17-35
VSB_HD float fresnel_conductor(float eta, float k, float cosi)
{
    const float ek = eta * eta + k * k;
    const float rs2 = (ek - 2.0f * eta * cosi + cosi * cosi) / (ek + 2.0f * eta * cosi + cosi * cosi);
    const float rp2 = (ek * cosi * cosi - 2.0f * eta * cosi + 1.0f) / (ek * cosi * cosi + 2.0f * eta * cosi + 1.0f);
    return (rs2 + rp2) / 2.0f;
}

// surface::sample -> generic_material::sample -- surface.h:50-92, detail/generic_material.inl:45-57:
// emissive pdf 1, direction untouched (emissive.inl:29-44); mirror = specular_reflection::sample_f
// (brdf.h:187-210); plastic / matte as before, times tex_color when the params carry one
VSB_HD V3 material_sample_ext(const Material& m, const Surface& S, V3 n, V3 wo, V3& wi, float& pdf, uint32_t& samp)
{
    if (m.kind == 2) { pdf = 1.0f; return v3(S.tex.x * m.cd[0] * m.kd, S.tex.y * m.cd[1] * m.kd, S.tex.z * m.cd[2] * m.kd); }
    if (m.kind == 3)
    {
        wi = vreflect(wo, n);
        pdf = 1.0f;
        const float cosi = fabsf(vdot(n, wo)), d = fabsf(vdot(n, wi));
        return v3(fresnel_conductor(m.ca[0], m.cd[0], cosi) * (m.cs[0] * m.ks) / d,
                  fresnel_conductor(m.ca[1], m.cd[1], cosi) * (m.cs[1] * m.ks) / d,
                  fresnel_conductor(m.ca[2], m.cd[2], cosi) * (m.cs[2] * m.ks) / d);
    }
    const V3 r = material_sample(m, n, wo, wi, pdf, samp);
    return v3(r.x * S.tex.x, r.y * S.tex.y, r.z * S.tex.z);
}

// ---- colour formats of the render target (pixel_access store / get, detail/color_conversion.h) ----
#if defined(__CUDACC__)
__device__ __forceinline__ int pixel_bytes(int format) { return format == 1 ? 4 : (format == 3 ? 3 : 16); }

// store(): RGBA -> RGB multiplies by alpha (color_conversion.h:231-247)
__device__ inline void store_pixel(int format, void* base, size_t index, float4 c)
{
    if (format == 0) { reinterpret_cast<float4*>(base)[index] = c; return; }
    if (format == 1)
    {
        reinterpret_cast<uchar4*>(base)[index] = make_uchar4((unsigned char)float_to_unorm8(c.x), (unsigned char)float_to_unorm8(c.y),
                                                             (unsigned char)float_to_unorm8(c.z), (unsigned char)float_to_unorm8(c.w));
        return;
    }
    const float r = rn_mul(c.x, c.w), g = rn_mul(c.y, c.w), b = rn_mul(c.z, c.w);
    if (format == 2)
    {
        float* q = reinterpret_cast<float*>(base) + index * 4;       // 16-byte pixels, 4th word untouched
        q[0] = r; q[1] = g; q[2] = b;
        return;
    }
    unsigned char* q = reinterpret_cast<unsigned char*>(base) + index * 3;
    q[0] = (unsigned char)float_to_unorm8(r); q[1] = (unsigned char)float_to_unorm8(g); q[2] = (unsigned char)float_to_unorm8(b);
}

// get(): RGB -> RGBA sets alpha 1 (color_conversion.h:249-265)
__device__ inline float4 load_pixel(int format, const void* base, size_t index)
{
    if (format == 0) return reinterpret_cast<const float4*>(base)[index];
    if (format == 1)
    {
        const uchar4 q = reinterpret_cast<const uchar4*>(base)[index];
        return make_float4(unorm8_to_float(q.x), unorm8_to_float(q.y), unorm8_to_float(q.z), unorm8_to_float(q.w));
    }
    if (format == 2)
    {
        const float* q = reinterpret_cast<const float*>(base) + index * 4;
        return make_float4(q[0], q[1], q[2], 1.0f);
    }
    const unsigned char* q = reinterpret_cast<const unsigned char*>(base) + index * 3;
    return make_float4(unorm8_to_float(q[0]), unorm8_to_float(q[1]), unorm8_to_float(q[2]), 1.0f);
}

// what a pixel reads back as after it was stored: get(store(c)) without touching memory (the SSAA
// accumulator of a formatted target lives in the lane state between sub-samples)
__device__ inline float4 roundtrip_pixel(int format, float4 c)
{
    if (format == 0) return c;
    if (format == 1)
        return make_float4(unorm8_to_float(float_to_unorm8(c.x)), unorm8_to_float(float_to_unorm8(c.y)),
                           unorm8_to_float(float_to_unorm8(c.z)), unorm8_to_float(float_to_unorm8(c.w)));
    const float r = rn_mul(c.x, c.w), g = rn_mul(c.y, c.w), b = rn_mul(c.z, c.w);
    if (format == 2) return make_float4(r, g, b, 1.0f);
    return make_float4(unorm8_to_float(float_to_unorm8(r)), unorm8_to_float(float_to_unorm8(g)), unorm8_to_float(float_to_unorm8(b)), 1.0f);
}
#endif

// ---- frame programs ------------------------------------------------------------------------------
//
// The built-in kernels as CONTINUATIONS around the persistent traversal loop (kernels.cuh,
// trace_kernel): begin() turns a pixel into its primary ray, done() receives the finished ray's hit
// record and either emits the next ray of that pixel (shadow ray, reflection, AO sample, path
// bounce) or writes the pixel and frees the lane.  Per-pixel state lives in shared memory
// (structure-of-arrays, one column per thread), not in registers, so the traversal loop keeps its
// register budget.  The arithmetic and the order of random draws are those of the reference
// kernels, statement for statement.

#if defined(__CUDACC__)

#define VSB_TILE 64
#ifndef VSB_FRAME_CHUNK
#define VSB_FRAME_CHUNK 32          // half an 8x8 pixel block (8x4) per warp reservation: short tail per launch
#endif

// a ray handed to the traversal loop by a program
struct RayReq
{
    float ox, oy, oz, dx, dy, dz;
    float max_t;
    bool  any;
};

struct FrameLane
{
    uint32_t* base;
    __device__ FrameLane(uint32_t* raw, unsigned tid) : base(raw + tid) {}
    __device__ float&    f(int k) { return reinterpret_cast<float*>(base)[k * 128]; }
    __device__ uint32_t& u(int k) { return base[k * 128]; }
    __device__ V3   ld(int k) { return v3(f(k), f(k + 1), f(k + 2)); }
    __device__ void st(int k, V3 a) { f(k) = a.x; f(k + 1) = a.y; f(k + 2) = a.z; }
};

// common words: destination index, sampler state, phase, pixel index, SSAA sub-sample index + accumulator
enum { W_DST = 0, W_SAMP = 1, W_PHASE = 2, W_PIX = 3, W_SUB = 4, W_ACC = 5, W_DEPTH = 9, W_DACC = 10 };   // 3.. only with EXT
enum { PH_PRIMARY = 0, PH_BOUNCE = 1, PH_SHADOW = 2 };

// EXT = true is the instantiation for the less common sched_params: SSAA pixel samplers (sub-sample
// accumulator in the lane state, resampling branch in finish()), the matrix camera and the depth
// target.  The common case (camera basis, uniform / jittered samplers, colour only) runs the leaner one.
// SURF = true adds what the later make_kernel_params overloads carry: get_surface() with bindings, colours
// and textures, and the generic material list.  Separate from EXT so that neither pays for the
// other's code: these programs are instruction-fetch sensitive.
template <int KERNEL, bool EXT, bool SURF = false>
struct FrameProgram
{
    FrameCtx ctx;
    int scissor[4];
    float4* color;                   // pixels of color_format (EXT; the lean instantiation is RGBA32F only)
    int color_format;
    float* depth;                    // EXT only; may be null
    int tile_rank, tile_nranks;      // tile_nranks == 0: single rank, row-major target
    int tiles_x;
    int in_place;                    // tile_nranks > 1: store into the whole row-major frame (e.g. peer memory)
    int tile_begin;                  // first rank-local tile of this launch (host targets are rendered in bands)
    int unit_tiles, unit_stride;     // a band = every unit_stride-th run of unit_tiles tiles, so bands cost alike (0: one launch)
    int n_local_tiles;
    float4* mirror;                  // compact tile buffers: the whole row-major frame finished pixels are ALSO stored into (or null)

    using Lane = FrameLane;
#ifndef VSB_FRAME_MIN_BLOCKS
#define VSB_FRAME_MIN_BLOCKS 6
#endif
    static constexpr int MIN_BLOCKS = VSB_FRAME_MIN_BLOCKS;     // resident CTAs per SM the register budget is set for
#ifndef VSB_FRAME_MIN_BLOCKS_Q
#define VSB_FRAME_MIN_BLOCKS_Q 0     // 0: per kernel (below); 5 .. 8: every frame program (development)
#endif
    // CTAs of 128 threads per SM for trace_kernel_q: 6 (80 registers) for the simple / Whitted programs, 7 (72 registers) for
    // AO / path tracing (gpurun_out/r2_ab7.log: 7 everywhere = cfg2 -4 %, cfg3 +3 %, cfg4 +1.5 %; 5 everywhere = -4 ... -6 %)
    static constexpr int MIN_BLOCKS_Q = VSB_FRAME_MIN_BLOCKS_Q ? VSB_FRAME_MIN_BLOCKS_Q : (KERNEL >= 2 ? 7 : 6);
    static constexpr bool GROUP_RING = VSB_GROUP_RING_ALL > 0 ? true : (VSB_GROUP_RING_ALL < 0 ? false : KERNEL >= 2);             // trace_kernel_q's ring holds (record range, mask, owner) groups: AO / path tracing
    static constexpr int K0 = EXT ? 11 : 3;      // first kernel-specific state word
    static constexpr int STATE_WORDS = K0 + (KERNEL == 1 ? (SURF ? 26 : 17) : (KERNEL == 0 ? 0 : 6));

    __device__ void closest(RayReq& r, V3 o, V3 d) const
    {
        r.ox = o.x; r.oy = o.y; r.oz = o.z; r.dx = d.x; r.dy = d.y; r.dz = d.z; r.max_t = FLT_MAX; r.any = false;
    }
    __device__ void occlusion(RayReq& r, V3 o, V3 d, float max_t) const
    {
        r.ox = o.x; r.oy = o.y; r.oz = o.z; r.dx = d.x; r.dy = d.y; r.dz = d.z; r.max_t = max_t; r.any = true;
    }

    // SSAA sample positions: detail/sched_common.h:222-295 (the 8x table's 0.1825 is the reference's)
    __device__ void ssaa_offset(int num, int k, float& dx, float& dy) const
    {
        if (num == 2) { dx = k == 0 ? -0.25f : 0.25f; dy = dx; return; }
        if (num == 4)
        {
            const float ox[4] = { -0.125f, 0.375f, 0.125f, -0.375f }, oy[4] = { -0.375f, -0.125f, 0.375f, 0.125f };
            dx = ox[k]; dy = oy[k]; return;
        }
        const float ox[8] = { -0.125f, 0.375f, -0.375f, 0.125f, -0.125f, 0.375f, -0.375f, 0.125f };
        const float oy[8] = { -0.4375f, -0.3125f, -0.1875f, -0.0625f, 0.0625f, 0.1825f, 0.3125f, 0.4375f };
        dx = ox[k]; dy = oy[k];
    }

    // per-sample kernel state + the primary ray of sub-sample k (k = 0 for the non-SSAA samplers)
    __device__ void start_sample(Lane& L, float fx, float fy, RayReq& req) const
    {
        L.u(W_PHASE) = PH_PRIMARY;
        if (KERNEL == 1)
        {
            L.st(K0 + 0, v3(0.0f, 0.0f, 0.0f));     // color
            L.f(K0 + 6) = 1.0f;                     // throughput
            L.u(K0 + 7) = 0u;                       // depth
        }
        if (KERNEL == 3)
        {
            L.st(K0 + 0, v3(1.0f, 1.0f, 1.0f));     // dst
            L.u(K0 + 3) = 0u;                       // bounce
            L.u(K0 + 4) = 0u;                       // result.hit
        }
        V3 o, d;
        if (EXT && ctx.camera_kind == 1) primary_ray_matrix(ctx, fx, fy, o, d);
        else primary_ray(ctx, fx, fy, o, d);
        if (EXT) L.f(W_DEPTH) = 1.0f;               // result.depth of a sample whose primary ray misses
        closest(req, o, d);
    }

    // result.hit / result.isect_pos of the built-in kernels are the PRIMARY hit (simple.inl:38,66;
    // whitted.inl:241; pathtracing.inl:68; ao/main.cpp:198-199): its depth is kept for finish()
    __device__ void primary_hit(Lane& L, V3 pos) const
    {
        if (EXT && depth) L.f(W_DEPTH) = depth_transform(ctx, pos);
    }

    // Pixel result of one kernel invocation.  uniform / jittered: store (detail/pixel_access.h:38-58);
    // jittered_blend: dst = src/frame_num + dst (1 - 1/frame_num) (detail/sched_common.h:516-528);
    // ssaa<N>: clear, then N x { dst = src/N + dst } (detail/sched_common.h:640-698) -- the
    // accumulator lives in the lane state and the next sub-sample's primary ray is emitted.
    // Returns true when the pixel continues with another ray (req).
    // finish() has ONE call site per program (done(), below).
    // It used to be inlined at every exit of a kernel: 6.6 k SASS instructions for the extended Whitted program and
    // stall_no_instruction on top of its profile (profiles/r1_surface_whitted_v0.txt).
    __device__ __forceinline__ bool finish(Lane& L, float r, float g, float b, float a, RayReq& req) const
    {
        return finish_body(L, r, g, b, a, req);
    }
    __device__ __forceinline__ bool finish_body(Lane& L, float r, float g, float b, float a, RayReq& req) const
    {
        const size_t dst = L.u(W_DST);
        float4 c = make_float4(r, g, b, a);
        const int ps = ctx.p.pixel_sampler;
        if (EXT && ps >= 3)
        {
            const int num = 2 << (ps - 3);
            const float alpha = rn_div(1.0f, (float)num);
            c.x = rn_add(rn_mul(c.x, alpha), L.f(W_ACC + 0)); c.y = rn_add(rn_mul(c.y, alpha), L.f(W_ACC + 1));
            c.z = rn_add(rn_mul(c.z, alpha), L.f(W_ACC + 2)); c.w = rn_add(rn_mul(c.w, alpha), L.f(W_ACC + 3));
            const uint32_t k = L.u(W_SUB) + 1u;
            const float z = rn_add(rn_mul(L.f(W_DEPTH), alpha), L.f(W_DACC));
            if ((int)k < num)
            {
                L.f(W_DACC) = z;
                if (color_format != 0) c = roundtrip_pixel(color_format, c);     // the target holds the stored value between blends
                L.f(W_ACC + 0) = c.x; L.f(W_ACC + 1) = c.y; L.f(W_ACC + 2) = c.z; L.f(W_ACC + 3) = c.w;
                L.u(W_SUB) = k;
                const uint32_t pixel = L.u(W_PIX);
                float dx, dy;
                ssaa_offset(num, (int)k, dx, dy);
                start_sample(L, rn_add((float)(pixel % (uint32_t)ctx.p.width), dx), rn_add((float)(pixel / (uint32_t)ctx.p.width), dy), req);
                return true;
            }
            if (depth) depth[dst] = z;
        }
        else if (ps == 2)
        {
            const float alpha = rn_div(1.0f, (float)ctx.p.frame_num);
            const float ia = rn_sub(1.0f, alpha);
            const float4 old = EXT ? load_pixel(color_format, color, dst) : color[dst];
            c.x = rn_add(rn_mul(c.x, alpha), rn_mul(old.x, ia));
            c.y = rn_add(rn_mul(c.y, alpha), rn_mul(old.y, ia));
            c.z = rn_add(rn_mul(c.z, alpha), rn_mul(old.z, ia));
            c.w = rn_add(rn_mul(c.w, alpha), rn_mul(old.w, ia));
            if (EXT && depth) depth[dst] = rn_add(rn_mul(L.f(W_DEPTH), alpha), rn_mul(depth[dst], ia));
        }
        else if (EXT && depth) depth[dst] = L.f(W_DEPTH);
        if (EXT) store_pixel(color_format, color, dst, c);
        else color[dst] = c;
        if (mirror)
        {
            // dst indexes the rank's compact tile buffer: back to the pixel's row-major position in the whole frame
            const uint32_t lt = (uint32_t)dst / (VSB_TILE * VSB_TILE), in = (uint32_t)dst % (VSB_TILE * VSB_TILE);
            const uint32_t tile = lt * (uint32_t)tile_nranks + (uint32_t)tile_rank;
            const uint32_t x = (tile % (uint32_t)tiles_x) * VSB_TILE + (in % VSB_TILE), y = (tile / (uint32_t)tiles_x) * VSB_TILE + (in / VSB_TILE);
            mirror[(size_t)y * (size_t)ctx.p.width + x] = c;
        }
        return false;
    }
    __device__ bool finish_bg(Lane& L, RayReq& req) const { return finish(L, ctx.p.bg[0], ctx.p.bg[1], ctx.p.bg[2], ctx.p.bg[3], req); }

    __device__ V3 ambient_term(const Material& m) const
    {
        if (SURF && m.kind >= 2) return v3(0.0f, 0.0f, 0.0f);     // emissive / mirror: detail/material/{emissive,mirror}.inl:14-18
        const float aw = ctx.p.ambient[3];
        return v3(m.ca[0] * m.ka * (ctx.p.ambient[0] * aw), m.ca[1] * m.ka * (ctx.p.ambient[1] * aw), m.ca[2] * m.ka * (ctx.p.ambient[2] * aw));
    }
    __device__ V3 face_normal(uint32_t prim_id) const
    {
        const float4 nn = __ldg(ctx.normals + prim_id);
        return v3(nn.x, nn.y, nn.z);
    }

    // item -> pixel: 64x64 tiles (tile k of the frame belongs to rank k % nranks), 8x8 blocks inside
    __device__ bool begin(Lane& L, unsigned long long item, RayReq& req) const
    {
        const FrameCtx& f = ctx;
        const int nr = tile_nranks > 1 ? tile_nranks : 1;
        const int rk = tile_nranks > 1 ? tile_rank : 0;
        int local_tile = (int)(item / (VSB_TILE * VSB_TILE));
        if (unit_stride > 1)
        {
            local_tile = tile_begin + (local_tile / unit_tiles) * (unit_stride * unit_tiles) + local_tile % unit_tiles;
            if (local_tile >= n_local_tiles) return false;
        }
        const int in_tile = (int)(item % (VSB_TILE * VSB_TILE));
        const int block = in_tile / 64, p = in_tile % 64;
        const int lx = (block % 8) * 8 + (p & 7), ly = (block / 8) * 8 + (p >> 3);
        const int tile = local_tile * nr + rk;
        const int x = (tile % tiles_x) * VSB_TILE + lx, y = (tile / tiles_x) * VSB_TILE + ly;
        const bool inside = x < f.p.width && y < f.p.height && x >= scissor[0] && y >= scissor[1] &&
                            x < scissor[0] + scissor[2] && y < scissor[1] + scissor[3];
        if (!inside) return false;
        const uint32_t pixel = (uint32_t)y * (uint32_t)f.p.width + (uint32_t)x;
        L.u(W_DST) = (tile_nranks > 1 && !in_place) ? (uint32_t)(local_tile * (VSB_TILE * VSB_TILE) + ly * VSB_TILE + lx) : pixel;
        if (EXT)
        {
            L.u(W_PIX) = pixel; L.u(W_SUB) = 0u;
            L.f(W_ACC + 0) = 0.0f; L.f(W_ACC + 1) = 0.0f; L.f(W_ACC + 2) = 0.0f;
            L.f(W_ACC + 3) = color_format >= 2 ? 1.0f : 0.0f;        // a cleared RGB pixel reads back with alpha 1
            L.f(W_DACC) = 0.0f;
        }
        uint32_t samp = sampler_init(pixel, f.p.frame_num, f.p.seed);
        float fx = (float)x, fy = (float)y;
        if (f.p.pixel_sampler == 1 || f.p.pixel_sampler == 2)
        {
            // jittered (detail/sched_common.h:197-217)
            const float jx = rn_sub(sampler_next(samp), 0.5f);
            const float jy = rn_sub(sampler_next(samp), 0.5f);
            fx = rn_add(fx, jx); fy = rn_add(fy, jy);
        }
        else if (EXT && f.p.pixel_sampler >= 3)
        {
            float dx, dy;
            ssaa_offset(2 << (f.p.pixel_sampler - 3), 0, dx, dy);
            fx = rn_add(fx, dx); fy = rn_add(fy, dy);
        }
        L.u(W_SAMP) = samp;
        start_sample(L, fx, fy, req);
        return true;
    }

    __device__ bool done(Lane& L, const HitReg& h, float ox, float oy, float oz, float dx, float dy, float dz, RayReq& req) const
    {
        const V3 o = v3(ox, oy, oz), d = v3(dx, dy, dz);
        // the kernels report a finished pixel through `px`; finish() -- store / blend / formats / SSAA resampling -- is
        // instantiated HERE only, however many exits a kernel has (these programs are instruction-fetch sensitive)
        float4 px;
        const bool more = KERNEL == 0 ? done_simple(L, h, o, d, req, px)
                        : KERNEL == 1 ? done_whitted(L, h, o, d, req, px)
                        : KERNEL == 2 ? done_ao(L, h, o, d, req, px) : done_pathtracing(L, h, o, d, req, px);
        if (more) return true;
        return finish(L, px.x, px.y, px.z, px.w, req);
    }
    __device__ __forceinline__ bool pixel(float4& px, float r, float g, float b, float a) const { px = make_float4(r, g, b, a); return false; }
    __device__ __forceinline__ bool pixel_bg(float4& px) const { px = make_float4(ctx.p.bg[0], ctx.p.bg[1], ctx.p.bg[2], ctx.p.bg[3]); return false; }

    // simple::kernel -- detail/simple.inl:25-75
    __device__ __forceinline__ bool done_simple(Lane& L, const HitReg& h, V3 o, V3 d, RayReq& req, float4& px) const
    {
        const FrameCtx& f = ctx;
        if (!h.hit) { return pixel_bg(px); }
        const V3 pos = vadd(o, vscale(d, h.t));
        primary_hit(L, pos);
        const Material m = f.materials[h.geom_id];
        V3 shaded = ambient_term(m);
        const V3 view = vneg(d);
        if (SURF)
        {
            const Surface S = surface_at(f, h.prim_id, h.geom_id, h.list_index, h.u, h.v, pos);
            const V3 n = vfaceforward(S.sn, view, S.gn);
            for (int i = 0; i < f.p.n_lights; ++i)
            {
                const PointLight l = f.lights[i];
                const V3 wi = vnormalize(vsub(ld3(l.pos), pos));
                const V3 c = shade_light_ext(m, S, l, pos, n, view, wi);
                shaded = v3(shaded.x + c.x, shaded.y + c.y, shaded.z + c.z);
            }
            return pixel(px, shaded.x, shaded.y, shaded.z, 1.0f);
        }
        V3 n = face_normal(h.prim_id);
        n = vfaceforward(n, view, n);
        for (int i = 0; i < f.p.n_lights; ++i)
        {
            const PointLight l = f.lights[i];
            const V3 wi = vnormalize(vsub(ld3(l.pos), pos));
            const V3 c = shade_light(m, l, pos, n, view, wi);
            shaded = v3(shaded.x + c.x, shaded.y + c.y, shaded.z + c.z);
        }
        return pixel(px, shaded.x, shaded.y, shaded.z, 1.0f);
    }

    // whitted::kernel -- detail/whitted.inl:227-323.  State: color 0..2, shaded 3..5, throughput 6,
    // depth 7, light 8, pos 9..11, view 12..14, prim_id 15, geom_id 16; SURF: the surface of the current hit,
    // evaluated once -- shading normal (faceforwarded) 17..19, tex_color 20..22, reflection direction 23..25
    __device__ __forceinline__ bool done_whitted(Lane& L, const HitReg& h, V3 o, V3 d, RayReq& req, float4& px) const
    {
        const FrameCtx& f = ctx;
        const uint32_t phase = L.u(W_PHASE);
        V3 pos, view, shaded;
        uint32_t prim_id, geom_id, light;
        if (phase == PH_SHADOW)
        {
            pos = L.ld(K0 + 9); view = L.ld(K0 + 12); shaded = L.ld(K0 + 3);
            prim_id = L.u(K0 + 15); geom_id = L.u(K0 + 16); light = L.u(K0 + 8);
            if (!h.hit)
            {
                const Material m = f.materials[geom_id];
                const PointLight l = f.lights[light];
                const V3 wi = vnormalize(vsub(ld3(l.pos), pos));
                V3 c;
                if (SURF)
                {
                    Surface S;
                    S.tex = L.ld(K0 + 20);
                    c = shade_light_ext(m, S, l, pos, L.ld(K0 + 17), view, wi);
                }
                else
                {
                    const V3 sn = face_normal(prim_id);
                    c = shade_light(m, l, pos, vfaceforward(sn, view, sn), view, wi);
                }
                shaded = v3(shaded.x + c.x, shaded.y + c.y, shaded.z + c.z);
            }
            ++light;
        }
        else
        {
            if (phase == PH_PRIMARY && !h.hit) { return pixel_bg(px); }
            if (EXT && phase == PH_PRIMARY) primary_hit(L, vadd(o, vscale(d, h.t)));   // detail/whitted.inl:238-242
            // while (hit && throughput > epsilon && depth++ < num_bounces)
            const float throughput = L.f(K0 + 6);
            bool go = h.hit && throughput > f.p.epsilon;
            if (go) { const uint32_t depth = L.u(K0 + 7); go = depth < f.p.num_bounces; L.u(K0 + 7) = depth + 1u; }
            if (!go) { const V3 c = L.ld(K0 + 0); return pixel(px, c.x, c.y, c.z, 1.0f); }
            pos = vadd(o, vscale(d, h.t));
            view = vneg(d);
            prim_id = h.prim_id; geom_id = h.geom_id; light = 0u;
            shaded = ambient_term(f.materials[geom_id]);
            L.st(K0 + 9, pos); L.st(K0 + 12, view); L.u(K0 + 15) = prim_id; L.u(K0 + 16) = geom_id;
            if (SURF)
            {
                const Surface S = surface_at(f, prim_id, geom_id, h.list_index, h.u, h.v, pos);
                L.st(K0 + 17, vfaceforward(S.sn, view, S.gn)); L.st(K0 + 20, S.tex);
                L.st(K0 + 23, vreflect(view, S.sn));          // specular_bounce(material, view_dir, surf.shading_normal)
            }
        }
        if ((int)light < f.p.n_lights)
        {
            const V3 lpos = ld3(f.lights[light].pos);
            const V3 wi = vnormalize(vsub(lpos, pos));
            occlusion(req, vadd(pos, vscale(wi, f.p.epsilon)), wi, vlength(vsub(pos, lpos)));
            L.st(K0 + 3, shaded); L.u(K0 + 8) = light; L.u(W_PHASE) = PH_SHADOW;
            return true;
        }
        const float throughput = L.f(K0 + 6);
        V3 color = L.ld(K0 + 0);
        color = v3(color.x + shaded.x * throughput, color.y + shaded.y * throughput, color.z + shaded.z * throughput);
        const float kr = SURF ? specular_kr(f.materials[geom_id])
                             : ((f.materials[geom_id].kind == 0) ? 0.1f : 0.0f);   // specular_bounce, detail/whitted.inl:66-122
        if (kr > 0.0f)
        {
            const V3 dir = SURF ? L.ld(K0 + 23) : vreflect(view, face_normal(prim_id));
            closest(req, vadd(pos, vscale(dir, f.p.epsilon)), dir);
            L.st(K0 + 0, color); L.f(K0 + 6) = rn_mul(throughput, kr); L.u(W_PHASE) = PH_BOUNCE;
            return true;
        }
        return pixel(px, color.x, color.y, color.z, 1.0f);    // throughput * 0 ends the loop
    }

    // AO kernel -- src/examples/ao/main.cpp:183-246.  State: pos 0..2, prim_id 3, clr 4, i 5
    __device__ __forceinline__ bool done_ao(Lane& L, const HitReg& h, V3 o, V3 d, RayReq& req, float4& px) const
    {
        const FrameCtx& f = ctx;
        V3 pos; uint32_t prim_id, i; float clr;
        if (L.u(W_PHASE) == PH_PRIMARY)
        {
            if (!h.hit) { return pixel_bg(px); }
            pos = vadd(o, vscale(d, h.t));
            primary_hit(L, pos);
            prim_id = h.prim_id; clr = 1.0f; i = 0u;
            L.st(K0 + 0, pos); L.u(K0 + 3) = prim_id; L.u(W_PHASE) = PH_SHADOW;
        }
        else
        {
            pos = L.ld(K0 + 0); prim_id = L.u(K0 + 3); clr = L.f(K0 + 4); i = L.u(K0 + 5);
            if (h.hit) clr = rn_sub(clr, 1.0f / (float)f.p.ao_samples);
            ++i;
        }
        if ((int)i < f.p.ao_samples)
        {
            const V3 w = face_normal(prim_id);
            V3 u, v;
            ortho_basis(u, v, w);
            uint32_t samp = L.u(W_SAMP);
            const float u2 = sampler_next(samp);   // right-to-left argument evaluation (see lambert_sample)
            const float u1 = sampler_next(samp);
            L.u(W_SAMP) = samp;
            const V3 dir = hemisphere_dir(u, v, w, cosine_sample_hemisphere(u1, u2));
            occlusion(req, vadd(pos, vscale(dir, 1E-3f)), dir, f.p.ao_radius);
            L.f(K0 + 4) = clr; L.u(K0 + 5) = i;
            return true;
        }
        return pixel(px, clr, clr, clr, 1.0f);
    }

    // pathtracing::kernel -- detail/pathtracing.inl:29-121.  State: dst 0..2, bounce 3, result.hit 4
    __device__ __forceinline__ bool done_pathtracing(Lane& L, const HitReg& h, V3 o, V3 d, RayReq& req, float4& px) const
    {
        const FrameCtx& f = ctx;
        // num_bounces == 0: the loop of the reference kernel never runs and every sample returns the background
        // (the primary ray traced for it is wasted; not a case worth a second copy of finish())
        if (f.p.num_bounces == 0u) return pixel_bg(px);
        V3 dst = L.ld(K0 + 0);
        uint32_t bounce = L.u(K0 + 3);
        bool result_hit = L.u(K0 + 4) != 0u;
        const float aw = f.p.ambient[3];
        bool finished = false;
        if (!h.hit)
        {
            dst = v3(dst.x * (f.p.ambient[0] * aw), dst.y * (f.p.ambient[1] * aw), dst.z * (f.p.ambient[2] * aw));
            finished = true;
        }
        else
        {
            if (bounce == 0u) { result_hit = true; primary_hit(L, vadd(o, vscale(d, h.t))); }
            const V3 view = vneg(d);
            const Material m = f.materials[h.geom_id];
            V3 refl = v3(0.0f, 0.0f, 0.0f);
            float pdf = 0.0f;
            uint32_t samp = L.u(W_SAMP);
            V3 n, src;
            if (SURF)
            {
                const Surface S = surface_at(f, h.prim_id, h.geom_id, h.list_index, h.u, h.v, vadd(o, vscale(d, h.t)));
                n = vfaceforward(S.sn, view, S.gn);
                src = material_sample_ext(m, S, n, view, refl, pdf, samp);
            }
            else
            {
                const V3 sn = face_normal(h.prim_id);
                n = vfaceforward(sn, view, sn);
                src = material_sample(m, n, view, refl, pdf, samp);
            }
            L.u(W_SAMP) = samp;
            if (pdf <= 0.0f) { dst = v3(0.0f, 0.0f, 0.0f); finished = true; }
            else if (SURF && m.kind == 2)
            {
                // has_emissive_material(surf): the path ends here with its throughput times ce * ls (detail/pathtracing.inl:93-104)
                dst = v3(dst.x * src.x, dst.y * src.y, dst.z * src.z);
                finished = true;
            }
            else
            {
                const float w = vdot(n, refl) / pdf;
                dst = v3(dst.x * (src.x * w), dst.y * (src.y * w), dst.z * (src.z * w));
                ++bounce;
                if (bounce >= f.p.num_bounces) { dst = v3(0.0f, 0.0f, 0.0f); finished = true; }   // still active after the last bounce
                else
                {
                    const V3 pos = vadd(o, vscale(d, h.t));
                    closest(req, vadd(pos, vscale(refl, f.p.epsilon)), refl);
                }
            }
        }
        if (finished) return result_hit ? pixel(px, dst.x, dst.y, dst.z, 1.0f) : pixel_bg(px);
        L.st(K0 + 0, dst); L.u(K0 + 3) = bounce; L.u(K0 + 4) = result_hit ? 1u : 0u;
        return true;
    }
};

#endif // __CUDACC__

} // namespace vsb
