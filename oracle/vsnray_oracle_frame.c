/*
 * vsnray_oracle_frame.c -- TEST INFRASTRUCTURE ONLY (see vsnray_oracle.h).
 *
 * Scalar restatement of the frame() path: per pixel ray generation, the built-in kernels
 * (simple / whitted / pathtracing) with the plastic / matte + point-light subset, the AO
 * kernel of src/examples/ao/main.cpp, and the store / blend rules.  Paths cited are relative
 * to /root/reference/include/visionaray/.  Colour arithmetic uses libm (tanf/powf), so frame
 * parity is a tolerance (1/255 per channel), not bit-exactness; the ray queries inside go
 * through vso_intersect_bvh and are bit-exact, and the sin / cos / pow that define SAMPLED
 * DIRECTIONS go through vso_raymath.h (correctly rounded, the same bits as the device code).
 */
#define _GNU_SOURCE 1
#include "vsnray_oracle.h"
#include "vso_raymath.h"

#include <float.h>
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#define VSO_PI      3.14159265358979323846f
#define VSO_TWO_PI  6.28318530717958647692f
#define VSO_INV_PI  0.31830988618379067154f

typedef struct { float x, y, z; } v3;

static inline v3 v3m(float x, float y, float z) { v3 r = { x, y, z }; return r; }
static inline v3 add(v3 a, v3 b) { return v3m(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 sub(v3 a, v3 b) { return v3m(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 mul(v3 a, v3 b) { return v3m(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline v3 scale(v3 a, float s) { return v3m(a.x * s, a.y * s, a.z * s); }
static inline v3 neg(v3 a) { return v3m(-a.x, -a.y, -a.z); }
static inline float dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline v3 cross(v3 u, v3 v)
{
    return v3m(u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x);
}
static inline v3 normalize(v3 v) { return scale(v, 1.0f / sqrtf(dot(v, v))); }
static inline float length(v3 v) { return sqrtf(dot(v, v)); }
static inline v3 ld3(const float* p) { return v3m(p[0], p[1], p[2]); }
static inline float fmax0(float x) { return 0.0f < x ? x : 0.0f; }   /* max(0, x) */

/* math/detail/vector.inl:685-688 */
static inline v3 reflect(v3 i, v3 n) { return sub(scale(n, 2.0f * dot(n, i)), i); }
/* math/detail/vector.inl:674-681 */
static inline v3 faceforward(v3 n, v3 i, v3 nref) { return dot(nref, i) < 0.0f ? neg(n) : n; }

typedef struct
{
    const vso_frame_params* p;
    const vso_bvh* b;
    const float* normals;
    const vso_surface* surf;              /* NULL: per-face normals, no colours / textures */
    const vso_material* mats;
    const vso_point_light* lights;
    float* rgba;
    int color_format;                     /* 0 PF_RGBA32F (rgba above), else `color` in that format */
    void* color;
    float eye[3], cu[3], cv[3], cw[3];
    int use_matrix;                       /* sched_params carries view / proj matrices */
    float view[16], inv_view[16], proj[16], inv_proj[16];
    float* depth;                         /* PF_DEPTH32F target or NULL */
} frame_ctx;

static void closest(const frame_ctx* f, v3 o, v3 d, vso_hit* h, vso_counters* c)
{
    vso_ray r = { { o.x, o.y, o.z }, 0.0f, { d.x, d.y, d.z }, 0.0f };
    vso_intersect_bvh(0, &r, f->b, FLT_MAX, h, c);
}

static int occluded(const frame_ctx* f, v3 o, v3 d, float max_t, vso_counters* c)
{
    vso_ray r = { { o.x, o.y, o.z }, 0.0f, { d.x, d.y, d.z }, 0.0f };
    vso_hit h;
    vso_intersect_bvh(1, &r, f->b, max_t, &h, c);
    return h.hit;
}

/* point_light::intensity -- detail/point_light.inl:11-27 */
static v3 light_intensity(const vso_point_light* l, v3 pos)
{
    float dist = length(sub(ld3(l->pos), pos));
    float att = 1.0f / (l->att_const + l->att_lin * dist + l->att_quad * dist * dist);
    if (l->kind == 1)
    {
        /* spot_light::intensity -- detail/spot_light.inl:14-38: (cl kl) att spot */
        v3 light_dir = sub(ld3(l->pos), pos);
        float spot = dot(ld3(l->spot_dir), normalize(neg(light_dir)));
        spot = spot > l->spot_cos_cutoff ? powf(spot, l->spot_exponent) : 0.0f;
        return scale(scale(scale(ld3(l->cl), l->kl), att), spot);
    }
    return scale(scale(ld3(l->cl), l->kl), att);
}

/* lambertian::f -- brdf.h:33-41 */
static v3 lambert_f(const vso_material* m) { return scale(scale(ld3(m->cd), m->kd), VSO_INV_PI); }

/* blinn::f (plastic's specular lobe, material.h:322-323) -- brdf.h:107-119 */
static v3 blinn_f(const vso_material* m, v3 n, v3 wo, v3 wi)
{
    v3 h = normalize(add(wo, wi));
    float hdotn = fmax0(dot(h, n));
    v3 spec = scale(ld3(m->cs), m->ks);
    float sat = dot(wi, h);
    sat = sat < 0.0f ? 0.0f : (sat > 1.0f ? 1.0f : sat);
    float sch = powf(1.0f - sat, 5.0f);
    v3 schlick = add(spec, scale(sub(v3m(1.0f, 1.0f, 1.0f), spec), sch));
    float nfactor = (m->spec_exp + 2.0f) / (8.0f * VSO_PI);
    return scale(scale(schlick, nfactor), powf(hdotn, m->spec_exp));
}

/* The surface a kernel shades: get_surface() -- get_surface.h:330-530, surface.h:24-96.
 * has_tex <=> the kernel_params carry colours / textures, i.e. the shade_record has a
 * tex_color member and the materials' tex_color overloads are the ones instantiated. */
typedef struct { v3 gn, sn, tex; int has_tex; const vso_material* m; } surf_t;

/* plastic::shade -- detail/material/plastic.inl:18-37 (+ cd() with tex_color :176-183);
 * matte::shade -- matte.inl:18-39 (:136-142); emissive::shade -- emissive.inl:22-27,76-90;
 * mirror::shade -- mirror.inl:22-27 (specular_reflection::f == 0, brdf.h:176-185) */
static v3 shade(const surf_t* s, const vso_point_light* l, v3 pos, v3 n, v3 wo, v3 wi)
{
    const vso_material* m = s->m;
    if (m->kind == 2)
    {
        v3 ce = scale(ld3(m->cd), m->kd);
        return s->has_tex ? scale(mul(s->tex, ld3(m->cd)), m->kd) : ce;
    }
    if (m->kind == 3) return v3m(0.0f, 0.0f, 0.0f);
    float ndotl = fmax0(dot(n, wi));
    v3 f = lambert_f(m);
    if (s->has_tex) f = mul(s->tex, f);
    if (m->kind == 0) f = add(f, blinn_f(m, n, wo, wi));
    return scale(mul(scale(f, VSO_PI), light_intensity(l, pos)), ndotl);
}

/* material.ambient(): plastic / matte ca * ka; emissive and mirror 0 (emissive.inl:14-18, mirror.inl:14-18) */
static v3 ambient_of(const vso_material* m)
{
    return m->kind <= 1 ? scale(ld3(m->ca), m->ka) : v3m(0.0f, 0.0f, 0.0f);
}

/* specular_bounce(): detail/whitted.inl:66-122 -- matte 0, mirror kr, everything else
 * (plastic, emissive) the fall-through 0.1 */
static float specular_kr(const vso_material* m)
{
    return m->kind == 1 ? 0.0f : (m->kind == 3 ? m->ks : 0.1f);
}

static v3 rgb_of_rgba(const float c[4]) { return v3m(c[0] * c[3], c[1] * c[3], c[2] * c[3]); }   /* from_rgba, detail/spectrum.inl:375-378 */


/* map_tex_coord -- texture/detail/filter/common.h:24-48 */
static float map_tex_coord(float coord, int texsize, int mode)
{
    float N = (float)texsize;
    switch (mode)
    {
    case 1:   /* Mirror */
        if ((((int)floorf(coord)) & 1) == 1) return (float)(texsize - 1) / (float)texsize - (coord - floorf(coord));
        return coord - floorf(coord);
    case 0:   /* Wrap */
        return coord - floorf(coord);
    default:  /* Clamp: clamp(coord, 0, 1 - 1/N) */
    {
        float hi = 1.0f - 1.0f / N;
        return coord < 0.0f ? 0.0f : (coord > hi ? hi : coord);   /* max(lo, min(x, hi)) */
    }
    }
}

static inline float lerp1(float a, float b, float x) { return (1.0f - x) * a + x * b; }   /* math/detail/math.h:459-464 */

/* spline weights -- texture/detail/filter/common.h:104-192.  The expressions mix float products with double
 * literals exactly as written there (a is a float; "a * a" alone is a float product, "0.5 * a" a double one). */
static float bs_w0(float a) { return (float)((1.0 / 6.0) * (-(a * a * a) + 3.0 * a * a - 3.0 * a + 1.0)); }
static float bs_w1(float a) { return (float)((1.0 / 6.0) * (3.0 * a * a * a - 6.0 * a * a + 4.0)); }
static float bs_w2(float a) { return (float)((1.0 / 6.0) * (-3.0 * a * a * a + 3.0 * a * a + 3.0 * a + 1.0)); }
static float bs_w3(float a) { return (float)((1.0 / 6.0) * (a * a * a)); }
static float cs_w0(float a) { return (float)(-0.5 * a * a * a + a * a - 0.5 * a); }
static float cs_w1(float a) { return (float)(1.5 * a * a * a - 2.5 * a * a + 1.0); }
static float cs_w2(float a) { return (float)(-1.5 * a * a * a + 2.0 * a * a + 0.5 * a); }
static float cs_w3(float a) { return (float)(0.5 * a * a * a - 0.5 * a * a); }

static void tex_point(const vso_texture* tex, int x, int y, float out[4])
{
    const uint8_t* e = (const uint8_t*)tex->data + 4 * ((size_t)y * (size_t)tex->width + (size_t)x);
    for (int k = 0; k < 4; ++k) out[k] = (float)e[k];
}

/* nearest -- filter/nearest.h:45-66; linear -- filter/linear.h:61-98: the result in the unnormalised integer
 * domain (ReturnT = vector<4, int>: truncated) */
static void tex_nearest_q(const vso_texture* tex, float s, float t, int q[4])
{
    const int w = tex->width, h = tex->height;
    float cs = map_tex_coord(s, w, tex->address[0]), ct = map_tex_coord(t, h, tex->address[1]);
    float e[4];
    tex_point(tex, (int)(cs * (float)w), (int)(ct * (float)h), e);
    for (int k = 0; k < 4; ++k) q[k] = (int)e[k];
}

static void tex_linear_q(const vso_texture* tex, float s, float t, int q[4])
{
    const int w = tex->width, h = tex->height;
    float s1 = map_tex_coord(s - 0.5f / (float)w, w, tex->address[0]), t1 = map_tex_coord(t - 0.5f / (float)h, h, tex->address[1]);
    float s2 = map_tex_coord(s + 0.5f / (float)w, w, tex->address[0]), t2 = map_tex_coord(t + 0.5f / (float)h, h, tex->address[1]);
    int lx = (int)(s1 * (float)w), ly = (int)(t1 * (float)h);
    int hx = (int)(s2 * (float)w), hy = (int)(t2 * (float)h);
    float e00[4], e10[4], e01[4], e11[4];
    tex_point(tex, lx, ly, e00); tex_point(tex, hx, ly, e10); tex_point(tex, lx, hy, e01); tex_point(tex, hx, hy, e11);
    float u = s1 * (float)w - (float)lx, v = t1 * (float)h - (float)ly;
    for (int k = 0; k < 4; ++k)
    {
        float p1 = lerp1(e00[k], e10[k], u);
        float p2 = lerp1(e01[k], e11[k], u);
        q[k] = (int)lerp1(p1, p2, v);
    }
}

/* cubic_opt (BSpline): four bilinear lookups at the g / h offsets -- filter/cubic_opt.h:58-95 */
static void tex_bspline_q(const vso_texture* tex, float s, float t, int q[4])
{
    const int w = tex->width, h = tex->height;
    float x = s * (float)w - 0.5f, floorx = floorf(x), fracx = x - floorf(x);
    float y = t * (float)h - 0.5f, floory = floorf(y), fracy = y - floorf(y);
    float tmp00 = bs_w1(fracx) / (bs_w0(fracx) + bs_w1(fracx));
    float h_00 = (floorx - 0.5f + tmp00) / w;
    float tmp10 = bs_w3(fracx) / (bs_w2(fracx) + bs_w3(fracx));
    float h_10 = (floorx + 1.5f + tmp10) / w;
    float tmp01 = bs_w1(fracy) / (bs_w0(fracy) + bs_w1(fracy));
    float h_01 = (floory - 0.5f + tmp01) / h;
    float tmp11 = bs_w3(fracy) / (bs_w2(fracy) + bs_w3(fracy));
    float h_11 = (floory + 1.5f + tmp11) / h;
    int a[4], b[4], c[4], d[4];
    tex_linear_q(tex, h_00, h_01, a); tex_linear_q(tex, h_10, h_01, b);
    tex_linear_q(tex, h_00, h_11, c); tex_linear_q(tex, h_10, h_11, d);
    float g0x = bs_w0(fracx) + bs_w1(fracx), g1x = bs_w2(fracx) + bs_w3(fracx);
    float g0y = bs_w0(fracy) + bs_w1(fracy), g1y = bs_w2(fracy) + bs_w3(fracy);
    for (int k = 0; k < 4; ++k)
    {
        float f_0 = g0x * (float)a[k] + g1x * (float)b[k];
        float f_1 = g0x * (float)c[k] + g1x * (float)d[k];
        q[k] = (int)(g0y * f_0 + g1y * f_1);
    }
}

/* cubic with the cardinal-spline weights (CardinalSpline): 4 x 4 texels -- filter/cubic.h:85-140 */
static void tex_cardinal_q(const vso_texture* tex, float s, float t, int q[4])
{
    const int w = tex->width, h = tex->height;
    const float off[4] = { -1.5f, -0.5f, 0.5f, 1.5f };
    float cx[4], cy[4]; int px[4], py[4];
    for (int i = 0; i < 4; ++i)
    {
        cx[i] = map_tex_coord(off[i] < 0 ? s - (-off[i]) / (float)w : s + off[i] / (float)w, w, tex->address[0]);
        cy[i] = map_tex_coord(off[i] < 0 ? t - (-off[i]) / (float)h : t + off[i] / (float)h, h, tex->address[1]);
        px[i] = (int)(cx[i] * (float)w); py[i] = (int)(cy[i] * (float)h);
    }
    float ux = cx[1] * (float)w - (float)px[1], uy = cy[1] * (float)h - (float)py[1];
    float wx[4] = { cs_w0(ux), cs_w1(ux), cs_w2(ux), cs_w3(ux) }, wy[4] = { cs_w0(uy), cs_w1(uy), cs_w2(uy), cs_w3(uy) };
    float f[4][4];
    for (int j = 0; j < 4; ++j)
    {
        float e0[4], e1[4], e2[4], e3[4];
        tex_point(tex, px[0], py[j], e0); tex_point(tex, px[1], py[j], e1); tex_point(tex, px[2], py[j], e2); tex_point(tex, px[3], py[j], e3);
        for (int k = 0; k < 4; ++k) f[j][k] = wx[0] * e0[k] + wx[1] * e1[k] + wx[2] * e2[k] + wx[3] * e3[k];
    }
    for (int k = 0; k < 4; ++k) q[k] = (int)(wy[0] * f[0][k] + wy[1] * f[1][k] + wy[2] * f[2][k] + wy[3] * f[3][k]);
}

/* tex2D for vector<4, unorm<8>> texels: filtering runs on the unnormalised integers, the result is
 * converted to vector<4, int> (truncation) and normalised once, float(uint32(q)) / double(255) -- a spline's
 * undershoot below 0 wraps through uint32 like in the reference
 * (texture/detail/sampler2d.h:81-113, math/detail/unorm.inl:26-31, filter.h:22-80) */
void vso_tex2d(const vso_texture* tex, float s, float t, float out[4])
{
    int q[4];
    switch (tex->filter)
    {
    case 1:  tex_linear_q(tex, s, t, q); break;
    case 2:  tex_bspline_q(tex, s, t, q); break;      /* BSpline */
    case 4:  tex_cardinal_q(tex, s, t, q); break;     /* CardinalSpline */
    default: tex_nearest_q(tex, s, t, q); break;      /* Nearest; BSplineInterpol (3) needs the prefilter, not restated */
    }
    for (int k = 0; k < 4; ++k) out[k] = (float)((double)(float)(uint32_t)q[k] / (double)255);
}

/* lerp(a, b, c, u, v) -- math/detail/math.h:466-475: (a (1 - (u + v)) + c v) + b u */
static v3 lerp3(v3 a, v3 b, v3 c, float u, float v)
{
    v3 s2 = scale(c, v), s3 = scale(b, u), s1 = scale(a, 1.0f - (u + v));
    return add(add(s1, s2), s3);
}

/* get_surface(hit_rec, params) for an index_bvh of triangles -- get_surface.h:330-530 */
/* get_normal(hr, primitive) -- the kernel_params carry no normals (kernels.h:94-140, 328-366):
 * get_surface.h:133-150 -> get_normal.h:96-101 triangle normalize(cross(e1, e2)); :113-118 sphere
 * (isect_pos - center) / radius; generic_primitive: the alternative's rule (detail/generic_primitive_get_normal.h) */
static v3 primitive_normal(const vso_bvh* b, uint32_t list_index, v3 pos)
{
    size_t pi = b->indices ? b->indices[list_index] : list_index;
    const void* rec; int tri;
    if (b->prim_kind == 0)      { rec = (const vso_triangle*)b->prims + pi; tri = 1; }
    else if (b->prim_kind == 1) { rec = (const vso_sphere*)b->prims + pi; tri = 0; }
    else
    {
        const vso_generic_primitive* g = (const vso_generic_primitive*)b->prims + pi;
        rec = g->storage; tri = g->which == 1;
    }
    if (tri)
    {
        const vso_triangle* t = (const vso_triangle*)rec;
        return normalize(cross(ld3(t->e1), ld3(t->e2)));
    }
    const vso_sphere* sp = (const vso_sphere*)rec;
    v3 d = sub(pos, ld3(sp->center));
    return v3m(d.x / sp->radius, d.y / sp->radius, d.z / sp->radius);
}

static void get_surface(const frame_ctx* f, const vso_hit* h, v3 pos, surf_t* S)
{
    const vso_surface* sf = f->surf;
    S->m = &f->mats[h->geom_id];
    S->has_tex = 0; S->tex = v3m(1.0f, 1.0f, 1.0f);
    if (!sf)
    {
        S->gn = S->sn = ld3(f->normals + 4 * (size_t)h->prim_id);   /* normals[hr.prim_id], get_normal.h:28-37 */
        return;
    }
    /* a generic_primitive: the visitor applies the alternative's rule (detail/generic_primitive_get_normal.h) -- a
     * sphere has no per-primitive normals under any binding (num_normals == 0) and takes get_normal(hr, sphere) */
    const vso_generic_primitive* gp = NULL;
    if (f->b->prim_kind == 2)
    {
        size_t gi = f->b->indices ? f->b->indices[h->prim_list_index] : h->prim_list_index;
        gp = (const vso_generic_primitive*)f->b->prims + gi;
    }
    if (gp && gp->which != 1) S->gn = S->sn = primitive_normal(f->b, h->prim_list_index, pos);
    else if (sf->normal_binding == 0) S->gn = S->sn = ld3(sf->normals + 4 * (size_t)h->prim_id);
    else if (sf->normal_binding == 2) S->gn = S->sn = primitive_normal(f->b, h->prim_list_index, pos);
    else
    {
        /* params.prims.begin[i].primitive(hr.primitive_list_index), get_surface.h:313-327; bvh.h:226-229 */
        const vso_bvh* b = f->b;
        size_t pi = b->indices ? b->indices[h->prim_list_index] : h->prim_list_index;
        const vso_triangle* tri = gp ? (const vso_triangle*)gp->storage : (const vso_triangle*)b->prims + pi;
        S->gn = normalize(cross(ld3(tri->e1), ld3(tri->e2)));
        const float* n = sf->normals + 12 * (size_t)h->prim_id;
        S->sn = normalize(lerp3(ld3(n), ld3(n + 4), ld3(n + 8), h->u, h->v));
    }
    if (sf->color_binding || sf->tex_coords)
    {
        S->has_tex = 1;
        v3 color = v3m(1.0f, 1.0f, 1.0f);
        if (sf->color_binding == 1) color = ld3(sf->colors + 4 * (size_t)h->prim_id);
        else if (sf->color_binding == 2)
        {
            const float* c = sf->colors + 12 * (size_t)h->prim_id;
            color = lerp3(ld3(c), ld3(c + 4), ld3(c + 8), h->u, h->v);
        }
        v3 tc = v3m(1.0f, 1.0f, 1.0f);
        if (sf->tex_coords && sf->textures && (int)h->geom_id < sf->n_textures)
        {
            const vso_texture* tex = &sf->textures[h->geom_id];
            if (tex->width > 0 && tex->height > 0)
            {
                const float* t = sf->tex_coords + 6 * (size_t)h->prim_id;
                /* lerp of vector<2,float>, get_tex_coord.h:24-40 */
                float w0 = 1.0f - (h->u + h->v);
                float s_ = (t[0] * w0 + t[4] * h->v) + t[2] * h->u;
                float t_ = (t[1] * w0 + t[5] * h->v) + t[3] * h->u;
                float q[4];
                vso_tex2d(tex, s_, t_, q);
                tc = v3m(q[0], q[1], q[2]);
            }
        }
        S->tex = sf->color_binding ? mul(color, tc) : tc;
    }
}

/* simple::kernel -- detail/simple.inl:25-75 */
static void kernel_simple(const frame_ctx* f, v3 o, v3 d, float out[4], vso_counters* c)
{
    const vso_frame_params* p = f->p;
    vso_hit h;
    closest(f, o, d, &h, c);
    if (!h.hit) { memcpy(out, p->bg, 16); return; }
    v3 pos = add(o, scale(d, h.t));
    surf_t S;
    get_surface(f, &h, pos, &S);
    v3 shaded = mul(ambient_of(S.m), rgb_of_rgba(p->ambient));
    v3 view = neg(d);
    v3 n = faceforward(S.sn, view, S.gn);
    for (int i = 0; i < p->n_lights; ++i)
    {
        v3 wi = normalize(sub(ld3(f->lights[i].pos), pos));
        shaded = add(shaded, shade(&S, &f->lights[i], pos, n, view, wi));
    }
    out[0] = shaded.x; out[1] = shaded.y; out[2] = shaded.z; out[3] = 1.0f;
}

/* whitted::kernel -- detail/whitted.inl:227-323; specular_bounce :79-110 (plastic: kr 0.1,
 * matte: kr 0) */
static void kernel_whitted(const frame_ctx* f, v3 o, v3 d, float out[4], vso_counters* c)
{
    const vso_frame_params* p = f->p;
    vso_hit h;
    closest(f, o, d, &h, c);
    if (!h.hit) { memcpy(out, p->bg, 16); return; }

    v3 color = v3m(0.0f, 0.0f, 0.0f);
    unsigned depth = 0;
    float throughput = 1.0f;
    while (h.hit && throughput > p->epsilon && depth++ < p->num_bounces)
    {
        v3 pos = add(o, scale(d, h.t));
        surf_t S;
        get_surface(f, &h, pos, &S);
        const vso_material* m = S.m;
        v3 sn = S.sn;
        v3 shaded = mul(ambient_of(m), rgb_of_rgba(p->ambient));
        v3 view = neg(d);
        v3 n = faceforward(sn, view, S.gn);
        for (int i = 0; i < p->n_lights; ++i)
        {
            v3 lpos = ld3(f->lights[i].pos);
            v3 wi = normalize(sub(lpos, pos));
            v3 so = add(pos, scale(wi, p->epsilon));
            float ld = length(sub(pos, lpos));
            if (!occluded(f, so, wi, ld, c))
                shaded = add(shaded, shade(&S, &f->lights[i], pos, n, view, wi));
        }
        color = add(color, scale(shaded, throughput));

        float kr = specular_kr(m);
        if (kr > 0.0f)
        {
            v3 dir = reflect(view, sn);
            o = add(pos, scale(dir, p->epsilon));
            d = dir;
            closest(f, o, d, &h, c);
        }
        throughput *= kr;
    }
    out[0] = color.x; out[1] = color.y; out[2] = color.z; out[3] = 1.0f;
}

/* make_orthonormal_basis -- math/detail/vector3.inl:357-367 */
static void ortho_basis(v3* u, v3* v, v3 w)
{
    if (fabsf(w.x) > fabsf(w.y)) *v = normalize(v3m(-w.z, 0.0f, w.x));
    else                         *v = normalize(v3m(0.0f, w.z, -w.y));
    *u = cross(*v, w);
}

/* cosine_sample_hemisphere -- sampling.h:61-71 */
static v3 cosine_sample_hemisphere(float u1, float u2)
{
    float r = sqrtf(u1);
    float theta = VSO_TWO_PI * u2;
    float sn, cs;
    vso_sincosf(theta, &sn, &cs);             /* std::cos / std::sin of the reference, correctly rounded (vso_raymath.h) */
    float x = r * cs;
    float y = r * sn;
    float z = sqrtf(fmax0(1.0f - u1));
    return v3m(x, y, z);
}

/* AO kernel -- src/examples/ao/main.cpp:183-246 (no faceforward; offset 1e-3 along dir) */
static void kernel_ao(const frame_ctx* f, v3 o, v3 d, uint32_t* samp, float out[4], vso_counters* c)
{
    const vso_frame_params* p = f->p;
    vso_hit h;
    closest(f, o, d, &h, c);
    if (!h.hit) { memcpy(out, p->bg, 16); return; }
    v3 pos = add(o, scale(d, h.t));
    v3 w = ld3(f->normals + 4 * (size_t)h.prim_id);
    v3 u, v;
    ortho_basis(&u, &v, w);
    float clr = 1.0f;
    float dec = 1.0f / (float)p->ao_samples;
    for (int i = 0; i < p->ao_samples; ++i)
    {
        /* cosine_sample_hemisphere(samp.next(), samp.next()): the argument evaluation order is
         * unspecified in C++; g++ (the compiler the reference is built with here) evaluates
         * right to left, so the FIRST draw is u2.  Pinned by tests/test_oracle_vs_ref.py. */
        float u2 = vso_sampler_next(samp);
        float u1 = vso_sampler_next(samp);
        v3 sp = cosine_sample_hemisphere(u1, u2);
        v3 dir = normalize(add(add(scale(u, sp.x), scale(v, sp.y)), scale(w, sp.z)));
        v3 ao_o = add(pos, scale(dir, 1E-3f));
        if (occluded(f, ao_o, dir, p->ao_radius, c)) clr = clr - dec;
    }
    out[0] = clr; out[1] = clr; out[2] = clr; out[3] = 1.0f;
}

/* lambertian::sample_f -- brdf.h:43-61 */
static v3 lambert_sample(const vso_material* m, v3 n, v3* wi, float* pdf, uint32_t* samp)
{
    v3 u, v;
    ortho_basis(&u, &v, n);
    float u2 = vso_sampler_next(samp);   /* right-to-left argument evaluation, see kernel_ao */
    float u1 = vso_sampler_next(samp);
    v3 sp = cosine_sample_hemisphere(u1, u2);
    *wi = normalize(add(add(scale(u, sp.x), scale(v, sp.y)), scale(n, sp.z)));
    *pdf = dot(n, *wi) * VSO_INV_PI;
    return lambert_f(m);
}

/* blinn::sample_f -- brdf.h:121-150 */
static v3 blinn_sample(const vso_material* m, v3 n, v3 wo, v3* wi, float* pdf, uint32_t* samp)
{
    float u1 = vso_sampler_next(samp);
    float u2 = vso_sampler_next(samp);
    float e = m->spec_exp;
    float costheta = vso_powf(u1, 1.0f / (e + 1.0f));   /* defines the sampled direction: vso_raymath.h */
    float sintheta = sqrtf(fmax0(1.0f - costheta * costheta));
    float phi = u2 * VSO_TWO_PI;
    v3 u, v;
    ortho_basis(&u, &v, n);
    float sp, cp;
    vso_sincosf(phi, &sp, &cp);
    v3 h = normalize(add(add(scale(u, sintheta * cp), scale(v, sintheta * sp)), scale(n, costheta)));
    *wi = reflect(wo, h);
    float vdoth = dot(wo, h);
    *pdf = ((e + 1.0f) * powf(costheta, e)) / (2.0f * VSO_PI * 4.0f * vdoth);
    return blinn_f(m, n, wo, *wi);
}

/* fresnel_reflectance(conductor_tag, eta, k, cosi) -- fresnel.h:17-35, per channel */
static float fresnel_conductor(float eta, float k, float cosi)
{
    float ek = eta * eta + k * k;
    float rs2 = (ek - 2.0f * eta * cosi + cosi * cosi) / (ek + 2.0f * eta * cosi + cosi * cosi);
    float rp2 = (ek * cosi * cosi - 2.0f * eta * cosi + 1.0f) / (ek * cosi * cosi + 2.0f * eta * cosi + 1.0f);
    return (rs2 + rp2) / 2.0f;
}

/* surface::sample -> material.sample(): plastic::sample_impl -- detail/material/plastic.inl:186-229
 * (x tex_color :48-63); matte -- matte.inl:147-155 (:50-65); emissive -- emissive.inl:29-44
 * (pdf 1, refl_dir untouched); mirror -- mirror.inl:29-41 -> specular_reflection::sample_f, brdf.h:187-210 */
static v3 material_sample(const surf_t* s, v3 n, v3 wo, v3* wi, float* pdf, uint32_t* samp)
{
    const vso_material* m = s->m;
    v3 r;
    if (m->kind == 2) { *pdf = 1.0f; return shade(s, NULL, v3m(0, 0, 0), n, wo, n); }
    if (m->kind == 3)
    {
        *wi = reflect(wo, n);
        *pdf = 1.0f;
        float cosi = fabsf(dot(n, wo)), d = fabsf(dot(n, *wi));
        return v3m(fresnel_conductor(m->ca[0], m->cd[0], cosi) * (m->cs[0] * m->ks) / d,
                   fresnel_conductor(m->ca[1], m->cd[1], cosi) * (m->cs[1] * m->ks) / d,
                   fresnel_conductor(m->ca[2], m->cd[2], cosi) * (m->cs[2] * m->ks) / d);
    }
    if (m->kind == 1) r = lambert_sample(m, n, wi, pdf, samp);
    else
    {
        float prob_diff = ((m->cd[0] + m->cd[1] + m->cd[2]) / 3.0f) * m->kd;
        float prob_spec = ((m->cs[0] + m->cs[1] + m->cs[2]) / 3.0f) * m->ks;
        if (prob_diff == 0.0f && prob_spec == 0.0f) { prob_diff = 0.5f; prob_spec = 0.5f; }
        prob_diff = prob_diff / (prob_diff + prob_spec);
        float u = vso_sampler_next(samp);
        if (u < prob_diff) r = lambert_sample(m, n, wi, pdf, samp);
        else               r = blinn_sample(m, n, wo, wi, pdf, samp);
    }
    return s->has_tex ? mul(s->tex, r) : r;
}

/* pathtracing::kernel -- detail/pathtracing.inl:29-121 */
static void kernel_pathtracing(const frame_ctx* f, v3 o, v3 d, uint32_t* samp, float out[4], vso_counters* c)
{
    const vso_frame_params* p = f->p;
    int active = 1, result_hit = 0;
    v3 dst = v3m(1.0f, 1.0f, 1.0f);
    for (unsigned bounce = 0; bounce < p->num_bounces; ++bounce)
    {
        vso_hit h;
        closest(f, o, d, &h, c);
        if (active && !h.hit) dst = mul(dst, rgb_of_rgba(p->ambient));
        active = active && h.hit;
        if (!active) break;
        if (bounce == 0) result_hit = h.hit;

        v3 view = neg(d);
        surf_t S;
        /* hit_rec.isect_pos is assigned only AFTER get_surface() here (detail/pathtracing.inl:74,107): a sphere's
         * normal (isect_pos - center) / radius would read an unset position in the reference, so binding 2 is
         * restricted to triangle scenes for this kernel (vsb_frame returns VSB_ERR_UNSUPPORTED otherwise) */
        get_surface(f, &h, add(o, scale(d, h.t)), &S);
        v3 n = faceforward(S.sn, view, S.gn);
        v3 refl = v3m(0.0f, 0.0f, 0.0f); float pdf = 0.0f;
        v3 src = material_sample(&S, n, view, &refl, &pdf, samp);
        int zero_pdf = pdf <= 0.0f;
        int emissive = S.m->kind == 2;                       /* has_emissive_material(surf), detail/surface.inl:14-38 */
        if (!emissive) src = scale(src, dot(n, refl) / pdf);
        if (!zero_pdf) dst = mul(dst, src);
        else           dst = v3m(0.0f, 0.0f, 0.0f);
        active = active && !emissive && !zero_pdf;
        if (!active) break;
        v3 pos = add(o, scale(d, h.t));
        o = add(pos, scale(refl, p->epsilon));
        d = refl;
    }
    if (active) dst = v3m(0.0f, 0.0f, 0.0f);
    if (result_hit) { out[0] = dst.x; out[1] = dst.y; out[2] = dst.z; out[3] = 1.0f; }
    else memcpy(out, p->bg, 16);
}

typedef struct { frame_ctx* f; size_t b, e; vso_counters c; } frame_job;

/* SSAA sample positions: detail/sched_common.h:222-295 (0.1825 in the 8x table is the reference's) */
static void ssaa_offset(int num, int k, float* dx, float* dy)
{
    static const float o4x[4] = { -0.125f, 0.375f, 0.125f, -0.375f }, o4y[4] = { -0.375f, -0.125f, 0.375f, 0.125f };
    static const float o8x[8] = { -0.125f, 0.375f, -0.375f, 0.125f, -0.125f, 0.375f, -0.375f, 0.125f };
    static const float o8y[8] = { -0.4375f, -0.3125f, -0.1875f, -0.0625f, 0.0625f, 0.1825f, 0.3125f, 0.4375f };
    if (num == 2) { *dx = k == 0 ? -0.25f : 0.25f; *dy = *dx; }
    else if (num == 4) { *dx = o4x[k]; *dy = o4y[k]; }
    else { *dx = o8x[k]; *dy = o8y[k]; }
}

/* ---- matrix camera ---------------------------------------------------------------------- */

#define M(m, r, c) ((m)[(c) * 4 + (r)])      /* matrix<4,4,T>::operator()(row, col), column-major */
static float det2(float m00, float m01, float m10, float m11) { return m00 * m11 - m10 * m01; }   /* math/detail/math.h det2 */

/* inverse(matrix<4,4,T>) -- math/detail/matrix4.inl:209-244 */
void vso_mat4_inverse(const float m[16], float out[16])
{
    float s0 = det2(M(m,0,0), M(m,0,1), M(m,1,0), M(m,1,1));
    float s1 = det2(M(m,0,0), M(m,0,2), M(m,1,0), M(m,1,2));
    float s2 = det2(M(m,0,0), M(m,0,3), M(m,1,0), M(m,1,3));
    float s3 = det2(M(m,0,1), M(m,0,2), M(m,1,1), M(m,1,2));
    float s4 = det2(M(m,0,1), M(m,0,3), M(m,1,1), M(m,1,3));
    float s5 = det2(M(m,0,2), M(m,0,3), M(m,1,2), M(m,1,3));
    float c5 = det2(M(m,2,2), M(m,2,3), M(m,3,2), M(m,3,3));
    float c4 = det2(M(m,2,1), M(m,2,3), M(m,3,1), M(m,3,3));
    float c3 = det2(M(m,2,1), M(m,2,2), M(m,3,1), M(m,3,2));
    float c2 = det2(M(m,2,0), M(m,2,3), M(m,3,0), M(m,3,3));
    float c1 = det2(M(m,2,0), M(m,2,2), M(m,3,0), M(m,3,2));
    float c0 = det2(M(m,2,0), M(m,2,1), M(m,3,0), M(m,3,1));
    float det = s0 * c5 - s1 * c4 + s2 * c3 + s3 * c2 - s4 * c1 + s5 * c0;
    /* the 16-scalar constructor takes the values column by column (matrix4.inl:28-39) */
    out[0]  = (+ M(m,1,1) * c5 - M(m,1,2) * c4 + M(m,1,3) * c3) / det;
    out[1]  = (- M(m,1,0) * c5 + M(m,1,2) * c2 + M(m,1,3) * c1) / det;
    out[2]  = (+ M(m,1,0) * c4 - M(m,1,1) * c2 + M(m,1,3) * c0) / det;
    out[3]  = (- M(m,1,0) * c3 + M(m,1,1) * c1 + M(m,1,2) * c0) / det;
    out[4]  = (- M(m,0,1) * c5 + M(m,0,2) * c4 - M(m,0,3) * c3) / det;
    out[5]  = (+ M(m,0,0) * c5 - M(m,0,2) * c2 + M(m,0,3) * c1) / det;
    out[6]  = (- M(m,0,0) * c4 + M(m,0,1) * c2 - M(m,0,3) * c0) / det;
    out[7]  = (+ M(m,0,0) * c3 - M(m,0,1) * c1 + M(m,0,2) * c0) / det;
    out[8]  = (+ M(m,3,1) * s5 - M(m,3,2) * s4 + M(m,3,3) * s3) / det;
    out[9]  = (- M(m,3,0) * s5 + M(m,3,2) * s2 - M(m,3,3) * s1) / det;
    out[10] = (+ M(m,3,0) * s4 - M(m,3,1) * s2 + M(m,3,3) * s0) / det;
    out[11] = (- M(m,3,0) * s3 + M(m,3,1) * s1 - M(m,3,2) * s0) / det;
    out[12] = (- M(m,2,1) * s5 + M(m,2,2) * s4 - M(m,2,3) * s3) / det;
    out[13] = (+ M(m,2,0) * s5 - M(m,2,2) * s2 + M(m,2,3) * s1) / det;
    out[14] = (- M(m,2,0) * s4 + M(m,2,1) * s2 - M(m,2,3) * s0) / det;
    out[15] = (+ M(m,2,0) * s3 - M(m,2,1) * s1 + M(m,2,2) * s0) / det;
}

/* matrix * vector<4> -- math/detail/matrix4.inl:171-181 */
static void mat4_mul(const float m[16], const float v[4], float out[4])
{
    for (int r = 0; r < 4; ++r) out[r] = M(m,r,0) * v[0] + M(m,r,1) * v[1] + M(m,r,2) * v[2] + M(m,r,3) * v[3];
}

/* make_primary_ray_impl, matrix overload -- detail/sched_common.h:152-178 */
void vso_primary_ray_matrix(float x, float y, int width, int height, const float inv_view[16],
                            const float inv_proj[16], vso_ray* out)
{
    float u = 2.0f * (x + 0.5f) / (float)width - 1.0f;
    float v = 2.0f * (y + 0.5f) / (float)height - 1.0f;
    float n4[4] = { u, v, -1.0f, 1.0f }, f4[4] = { u, v, 1.0f, 1.0f }, t[4], o[4], d[4];
    mat4_mul(inv_proj, n4, t); mat4_mul(inv_view, t, o);
    mat4_mul(inv_proj, f4, t); mat4_mul(inv_view, t, d);
    v3 ori = v3m(o[0] / o[3], o[1] / o[3], o[2] / o[3]);
    v3 dir = normalize(sub(v3m(d[0] / d[3], d[1] / d[3], d[2] / d[3]), ori));
    memset(out, 0, sizeof(*out));
    out->ori[0] = ori.x; out->ori[1] = ori.y; out->ori[2] = ori.z;
    out->dir[0] = dir.x; out->dir[1] = dir.y; out->dir[2] = dir.z;
}

/* depth_transform -- detail/sched_common.h:337-354 */
float vso_depth_transform(const float pos[3], const float view[16], const float proj[16])
{
    float p4[4] = { pos[0], pos[1], pos[2], 1.0f }, t[4], q[4];
    mat4_mul(view, p4, t); mat4_mul(proj, t, q);
    return (q[2] / q[3] + 1.0f) * 0.5f;
}
#undef M

/* One kernel invocation for the sample at (px, py).  *depth (matrix camera only) receives
 * result.depth = hit ? depth_transform(isect_pos) : 1 (detail/sched_common.h:425): every built-in
 * kernel reports the PRIMARY hit and its position (simple.inl:38,66; whitted.inl:241;
 * pathtracing.inl:68; ao/main.cpp:198-199), so it is evaluated from the primary ray alone. */
static void run_kernel(const frame_ctx* f, float px, float py, uint32_t* samp, float col[4], float* depth, vso_counters* c)
{
    const vso_frame_params* p = f->p;
    vso_ray r;
    if (f->use_matrix) vso_primary_ray_matrix(px, py, p->width, p->height, f->inv_view, f->inv_proj, &r);
    else vso_primary_ray(px, py, p->width, p->height, f->eye, f->cu, f->cv, f->cw, &r);
    v3 o = ld3(r.ori), d = ld3(r.dir);
    if (depth)
    {
        vso_hit h; vso_counters unused = { 0, 0, 0, 0 };
        *depth = 1.0f;
        if (!(p->kernel == 3 && p->num_bounces == 0u))
        {
            closest(f, o, d, &h, &unused);
            if (h.hit)
            {
                v3 pos = add(o, scale(d, h.t));
                float pp[3] = { pos.x, pos.y, pos.z };
                *depth = vso_depth_transform(pp, f->view, f->proj);
            }
        }
    }
    switch (p->kernel)
    {
    case 0: kernel_simple(f, o, d, col, c); break;
    case 1: kernel_whitted(f, o, d, col, c); break;
    case 2: kernel_ao(f, o, d, samp, col, c); break;
    default: kernel_pathtracing(f, o, d, samp, col, c); break;
    }
}

/* float_to_unorm<8> / unorm_to_float<8> -- math/detail/unorm.inl:18-31 */
static uint8_t to_unorm8(float f)
{
    f = f < 0.0f ? 0.0f : (f > 1.0f ? 1.0f : f);      /* clamp (color_conversion.h:224), saturate */
    return (uint8_t)(uint32_t)(f * (double)255);
}
static float from_unorm8(uint8_t u) { return (float)((float)u / (double)255); }

static const int vso_bpp[4] = { 16, 4, 16, 3 };   /* pixel_traits<PF_RGB32F>::type is vector<3,float>: 16 bytes, last word unused */

/* pixel_access::store -> convert(): detail/color_conversion.h:213-247 (RGBA -> RGB multiplies by alpha) */
void vso_store_pixel(int format, const float c[4], void* dst)
{
    if (format == 0) memcpy(dst, c, 16);
    else if (format == 1) { uint8_t* q = (uint8_t*)dst; for (int k = 0; k < 4; ++k) q[k] = to_unorm8(c[k]); }
    else
    {
        float rgb[3] = { c[0] * c[3], c[1] * c[3], c[2] * c[3] };
        if (format == 2) memcpy(dst, rgb, 12);
        else { uint8_t* q = (uint8_t*)dst; for (int k = 0; k < 3; ++k) q[k] = to_unorm8(rgb[k]); }
    }
}

/* pixel_access::get -> convert(): RGB -> RGBA sets alpha 1 (detail/color_conversion.h:249-265) */
void vso_get_pixel(int format, const void* src, float c[4])
{
    if (format == 0) memcpy(c, src, 16);
    else if (format == 1) { const uint8_t* q = (const uint8_t*)src; for (int k = 0; k < 4; ++k) c[k] = from_unorm8(q[k]); }
    else if (format == 2) { memcpy(c, src, 12); c[3] = 1.0f; }
    else { const uint8_t* q = (const uint8_t*)src; for (int k = 0; k < 3; ++k) c[k] = from_unorm8(q[k]); c[3] = from_unorm8(to_unorm8(1.0f)); }
}

/* store / blend of one kernel result into a formatted target: detail/pixel_access.h:1153-1175 */
static void blend_pixel(int format, void* dst, const float col[4], float sfactor, float dfactor)
{
    float d[4];
    vso_get_pixel(format, dst, d);
    for (int k = 0; k < 4; ++k) d[k] = col[k] * sfactor + d[k] * dfactor;
    vso_store_pixel(format, d, dst);
}

static void frame_range_format(frame_ctx* f, size_t b, size_t e, vso_counters* c)
{
    const vso_frame_params* p = f->p;
    const int fmt = f->color_format;
    size_t w = (size_t)p->width;
    for (size_t i = b; i < e; ++i)
    {
        unsigned x = (unsigned)(i % w), y = (unsigned)(i / w);
        uint32_t samp = vso_sampler_init((uint32_t)i, p->frame_num, p->seed);
        float px = (float)x, py = (float)y;
        char* dst = (char*)f->color + i * (size_t)vso_bpp[fmt];
        float col[4];
        if (p->pixel_sampler >= 3)
        {
            const int num = 2 << (p->pixel_sampler - 3);
            const float alpha = 1.0f / (float)num;
            const float zero[4] = { 0.0f, 0.0f, 0.0f, 0.0f };
            vso_store_pixel(fmt, zero, dst);
            for (int s = 0; s < num; ++s)
            {
                float dx, dy;
                ssaa_offset(num, s, &dx, &dy);
                run_kernel(f, px + dx, py + dy, &samp, col, NULL, c);
                blend_pixel(fmt, dst, col, alpha, 1.0f);
            }
            continue;
        }
        if (p->pixel_sampler != 0)
        {
            float jx = vso_sampler_next(&samp) - 0.5f;
            float jy = vso_sampler_next(&samp) - 0.5f;
            px += jx; py += jy;
        }
        run_kernel(f, px, py, &samp, col, NULL, c);
        if (p->pixel_sampler == 2)
        {
            float alpha = 1.0f / (float)p->frame_num;
            blend_pixel(fmt, dst, col, alpha, 1.0f - alpha);
        }
        else vso_store_pixel(fmt, col, dst);
    }
}

static void frame_range(frame_ctx* f, size_t b, size_t e, vso_counters* c)
{
    if (f->color_format != 0) { frame_range_format(f, b, e, c); return; }
    const vso_frame_params* p = f->p;
    size_t w = (size_t)p->width;
    for (size_t i = b; i < e; ++i)
    {
        unsigned x = (unsigned)(i % w), y = (unsigned)(i / w);
        uint32_t samp = vso_sampler_init((uint32_t)i, p->frame_num, p->seed);
        float px = (float)x, py = (float)y;
        float* dst = f->rgba + i * 4;
        float* zdst = f->depth ? f->depth + i : NULL;
        float col[4], z = 0.0f;
        float* zp = zdst ? &z : NULL;
        if (p->pixel_sampler >= 3)
        {
            /* ssaa_type<N>: clear, then N x blend(src / N + dst): detail/sched_common.h:640-698 */
            const int num = 2 << (p->pixel_sampler - 3);
            const float alpha = 1.0f / (float)num;
            for (int k = 0; k < 4; ++k) dst[k] = 0.0f;
            if (zdst) *zdst = 0.0f;
            for (int s = 0; s < num; ++s)
            {
                float dx, dy;
                ssaa_offset(num, s, &dx, &dy);
                run_kernel(f, px + dx, py + dy, &samp, col, zp, c);
                for (int k = 0; k < 4; ++k) dst[k] = col[k] * alpha + dst[k] * 1.0f;
                if (zdst) *zdst = z * alpha + *zdst * 1.0f;
            }
            continue;
        }
        if (p->pixel_sampler != 0)
        {
            /* jittered: detail/sched_common.h:197-217 */
            float jx = vso_sampler_next(&samp) - 0.5f;
            float jy = vso_sampler_next(&samp) - 0.5f;
            px += jx; py += jy;
        }
        run_kernel(f, px, py, &samp, col, zp, c);
        if (p->pixel_sampler == 2)
        {
            /* jittered_blend: detail/sched_common.h:516-528 (with depth: :559-575) */
            float alpha = 1.0f / (float)p->frame_num;
            for (int k = 0; k < 4; ++k) dst[k] = col[k] * alpha + dst[k] * (1.0f - alpha);
            if (zdst) *zdst = z * alpha + *zdst * (1.0f - alpha);
        }
        else
        {
            /* uniform / jittered store: detail/pixel_access.h:38-58, :608-658 */
            for (int k = 0; k < 4; ++k) dst[k] = col[k];
            if (zdst) *zdst = z;
        }
    }
}

typedef struct { frame_ctx* f; size_t n; size_t* next; pthread_mutex_t* mu; vso_counters c; } fworker;

static void* fworker_main(void* q)
{
    fworker* w = (fworker*)q;
    for (;;)
    {
        size_t b;
        pthread_mutex_lock(w->mu);
        b = *w->next; *w->next += 1024;
        pthread_mutex_unlock(w->mu);
        if (b >= w->n) break;
        size_t e = b + 1024 < w->n ? b + 1024 : w->n;
        frame_range(w->f, b, e, &w->c);
    }
    return NULL;
}

static void frame_run(frame_ctx* fp, vso_counters* counters, int nthreads);

void vso_frame(const vso_frame_params* p, const vso_bvh* b, const float* normals16,
               const vso_material* mats, const vso_point_light* lights, float* rgba,
               vso_counters* counters, int nthreads)
{
    frame_ctx f;
    memset(&f, 0, sizeof(f));
    f.p = p; f.b = b; f.normals = normals16; f.mats = mats; f.lights = lights; f.rgba = rgba;
    vso_camera_basis(p, f.eye, f.cu, f.cv, f.cw);
    frame_run(&f, counters, nthreads);
}

void vso_frame_format(const vso_frame_params* p, const vso_bvh* b, const float* normals16,
                      const vso_material* mats, const vso_point_light* lights, void* color, int format,
                      vso_counters* counters, int nthreads)
{
    frame_ctx f;
    memset(&f, 0, sizeof(f));
    f.p = p; f.b = b; f.normals = normals16; f.mats = mats; f.lights = lights;
    f.rgba = (float*)color; f.color = color; f.color_format = format;
    vso_camera_basis(p, f.eye, f.cu, f.cv, f.cw);
    frame_run(&f, counters, nthreads);
}

void vso_frame_surface(const vso_frame_params* p, const vso_bvh* b, const vso_surface* surf,
                       const vso_material* mats, const vso_point_light* lights, float* rgba,
                       vso_counters* counters, int nthreads)
{
    frame_ctx f;
    memset(&f, 0, sizeof(f));
    f.p = p; f.b = b; f.surf = surf; f.normals = surf ? surf->normals : NULL; f.mats = mats; f.lights = lights; f.rgba = rgba;
    vso_camera_basis(p, f.eye, f.cu, f.cv, f.cw);
    frame_run(&f, counters, nthreads);
}

/* the scheduler computes both inverses on the host each frame (detail/tiled_sched.inl:241-242,
 * detail/cuda_sched.inl:231-234) */
void vso_frame_matrix(const vso_frame_params* p, const float view[16], const float proj[16],
                      const vso_bvh* b, const float* normals16, const vso_material* mats,
                      const vso_point_light* lights, float* rgba, float* depth,
                      vso_counters* counters, int nthreads)
{
    frame_ctx f;
    memset(&f, 0, sizeof(f));
    f.p = p; f.b = b; f.normals = normals16; f.mats = mats; f.lights = lights; f.rgba = rgba;
    f.use_matrix = 1; f.depth = depth;
    memcpy(f.view, view, 64); memcpy(f.proj, proj, 64);
    vso_mat4_inverse(view, f.inv_view); vso_mat4_inverse(proj, f.inv_proj);
    frame_run(&f, counters, nthreads);
}

static void frame_run(frame_ctx* fp, vso_counters* counters, int nthreads)
{
    frame_ctx f = *fp;
    const vso_frame_params* p = f.p;
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    size_t n = (size_t)p->width * (size_t)p->height, next = 0;
    pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;
    fworker* ws = (fworker*)calloc((size_t)nthreads, sizeof(fworker));
    pthread_t* ts = (pthread_t*)calloc((size_t)nthreads, sizeof(pthread_t));
    for (int i = 0; i < nthreads; ++i) { ws[i].f = &f; ws[i].n = n; ws[i].next = &next; ws[i].mu = &mu; }
    if (nthreads == 1) fworker_main(&ws[0]);
    else
    {
        for (int i = 0; i < nthreads; ++i) pthread_create(&ts[i], NULL, fworker_main, &ws[i]);
        for (int i = 0; i < nthreads; ++i) pthread_join(ts[i], NULL);
    }
    if (counters)
        for (int i = 0; i < nthreads; ++i)
        {
            counters->n_inner += ws[i].c.n_inner; counters->n_prim += ws[i].c.n_prim;
            counters->n_rays += ws[i].c.n_rays;   counters->n_hits += ws[i].c.n_hits;
        }
    free(ws); free(ts);
}

/* vso_raymath.h through the library, for tests/test_raymath.py */
void vso_ray_sincosf(const float* x, size_t n, float* s, float* c)
{
    for (size_t i = 0; i < n; ++i) vso_sincosf(x[i], s + i, c + i);
}
void vso_ray_powf(const float* x, const float* y, size_t n, float* out)
{
    for (size_t i = 0; i < n; ++i) out[i] = vso_powf(x[i], y[i]);
}
