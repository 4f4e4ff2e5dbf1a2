// watertight.cuh -- opt-in traversal mode VSB_TRAVERSE_WATERTIGHT: the hit / miss DECISION of a (ray, triangle) pair
// comes from the watertight test of Woop, Benthin, Wald ("Watertight Ray/Triangle Intersection", JCGT 2(1), 2013:
// shear the triangle into ray space, three edge functions with a double-precision fallback when one is exactly 0, so
// that a ray cannot slip between two triangles that share an edge); t / u / v of an accepted hit are reported in the
// Moeller-Trumbore form of the reference (math/intersect.h:122-179), so records agree bit for bit with the default
// mode wherever both tests accept.  The default stays the reference's own test: its misses on shared edges are part
// of "results identical to the reference".  The boxes need nothing extra: the quantised BVH8 boxes contain their
// triangles and the slab test is inflated (bvh8.cuh), so a ray that touches a triangle's edge reaches its leaf.
//
// Limit of the data layout: the reference stores a triangle as (v1, e1, e2); the shared vertices of two neighbours
// are reconstructed as v1 + e1 / v1 + e2 with one rounding each and may differ in the last bit when e = v2 - v1 was
// not exact.  The mode is watertight with respect to the RECONSTRUCTED vertices (exactly so for meshes whose edge
// vectors are exact differences, e.g. dyadic coordinates; tests/test_watertight.py).
#pragma once

#include "bvh8_traverse.cuh"

namespace vsb
{

struct WtRay
{
    int   kx, ky, kz;
    float sx, sy, sz;
};

VSB_HD float wt_get(float x, float y, float z, int k) { return k == 0 ? x : (k == 1 ? y : z); }

VSB_HD void wt_ray_init(WtRay& w, float dx, float dy, float dz)
{
    const float ax = fabsf(dx), ay = fabsf(dy), az = fabsf(dz);
    w.kz = ax > ay ? (ax > az ? 0 : 2) : (ay > az ? 1 : 2);
    w.kx = w.kz == 2 ? 0 : w.kz + 1;
    w.ky = w.kx == 2 ? 0 : w.kx + 1;
    const float dk = wt_get(dx, dy, dz, w.kz);
    if (dk < 0.0f) { const int t = w.kx; w.kx = w.ky; w.ky = t; }       // preserve the winding
    w.sx = rn_div(wt_get(dx, dy, dz, w.kx), dk);
    w.sy = rn_div(wt_get(dx, dy, dz, w.ky), dk);
    w.sz = rn_div(1.0f, dk);
}

// the decision: does the ray's LINE pass through the closed triangle (v1, v1 + e1, v1 + e2)?  Two-sided, like the reference.
VSB_HD bool wt_hits(const WtRay& w, float ox, float oy, float oz, float v1x, float v1y, float v1z,
                    float e1x, float e1y, float e1z, float e2x, float e2y, float e2z)
{
    const float p2x = rn_add(v1x, e1x), p2y = rn_add(v1y, e1y), p2z = rn_add(v1z, e1z);
    const float p3x = rn_add(v1x, e2x), p3y = rn_add(v1y, e2y), p3z = rn_add(v1z, e2z);
    const float ax_ = rn_sub(v1x, ox), ay_ = rn_sub(v1y, oy), az_ = rn_sub(v1z, oz);
    const float bx_ = rn_sub(p2x, ox), by_ = rn_sub(p2y, oy), bz_ = rn_sub(p2z, oz);
    const float cx_ = rn_sub(p3x, ox), cy_ = rn_sub(p3y, oy), cz_ = rn_sub(p3z, oz);
    const float akz = wt_get(ax_, ay_, az_, w.kz), bkz = wt_get(bx_, by_, bz_, w.kz), ckz = wt_get(cx_, cy_, cz_, w.kz);
    const float Ax = rn_sub(wt_get(ax_, ay_, az_, w.kx), rn_mul(w.sx, akz)), Ay = rn_sub(wt_get(ax_, ay_, az_, w.ky), rn_mul(w.sy, akz));
    const float Bx = rn_sub(wt_get(bx_, by_, bz_, w.kx), rn_mul(w.sx, bkz)), By = rn_sub(wt_get(bx_, by_, bz_, w.ky), rn_mul(w.sy, bkz));
    const float Cx = rn_sub(wt_get(cx_, cy_, cz_, w.kx), rn_mul(w.sx, ckz)), Cy = rn_sub(wt_get(cx_, cy_, cz_, w.ky), rn_mul(w.sy, ckz));
    float U = rn_sub(rn_mul(Cx, By), rn_mul(Cy, Bx));
    float V = rn_sub(rn_mul(Ax, Cy), rn_mul(Ay, Cx));
    float W = rn_sub(rn_mul(Bx, Ay), rn_mul(By, Ax));
    if (U == 0.0f || V == 0.0f || W == 0.0f)
    {
        // exactly on an edge in float: decide in double (products of floats are exact there, one rounding in the difference)
        const double CxBy = (double)Cx * (double)By, CyBx = (double)Cy * (double)Bx;
        const double AxCy = (double)Ax * (double)Cy, AyCx = (double)Ay * (double)Cx;
        const double BxAy = (double)Bx * (double)Ay, ByAx = (double)By * (double)Ax;
        U = (float)(CxBy - CyBx); V = (float)(AxCy - AyCx); W = (float)(BxAy - ByAx);
    }
    if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return false;
    const float det = rn_add(rn_add(U, V), W);
    return det != 0.0f;
}

// t, u, v in the reference's Moeller-Trumbore form (isect_triangle's arithmetic) without its own inside test
VSB_HD bool mt_tuv(float ox, float oy, float oz, float dx, float dy, float dz, float v1x, float v1y, float v1z,
                   float e1x, float e1y, float e1z, float e2x, float e2y, float e2z, float& t, float& u, float& v)
{
    const float s1x = rn_sub(rn_mul(dy, e2z), rn_mul(dz, e2y));
    const float s1y = rn_sub(rn_mul(dz, e2x), rn_mul(dx, e2z));
    const float s1z = rn_sub(rn_mul(dx, e2y), rn_mul(dy, e2x));
    const float div = rn_dot(s1x, s1y, s1z, e1x, e1y, e1z);
    if (!(div != 0.0f)) return false;
    const float inv_div = rn_rcp(div);
    const float ddx = rn_sub(ox, v1x), ddy = rn_sub(oy, v1y), ddz = rn_sub(oz, v1z);
    u = rn_mul(rn_dot(ddx, ddy, ddz, s1x, s1y, s1z), inv_div);
    const float s2x = rn_sub(rn_mul(ddy, e1z), rn_mul(ddz, e1y));
    const float s2y = rn_sub(rn_mul(ddz, e1x), rn_mul(ddx, e1z));
    const float s2z = rn_sub(rn_mul(ddx, e1y), rn_mul(ddy, e1x));
    v = rn_mul(rn_dot(dx, dy, dz, s2x, s2y, s2z), inv_div);
    t = rn_mul(rn_dot(e2x, e2y, e2z, s2x, s2y, s2z), inv_div);
    return true;
}

// closest_hit / any_hit of one ray over one BVH8 with the watertight decision (triangle scenes)
template <bool ANY>
VSB_HD void trace_wide_watertight(const WideScene& s, const RayCtx& r, HitReg& best, float& limit)
{
    WtRay w;
    wt_ray_init(w, r.dx, r.dy, r.dz);
    uint32_t sb[VSB_STACK_SIZE], sm[VSB_STACK_SIZE];
    int sp = 0;
    uint32_t ng_base = 0u, ng_mask = 0x80000000u, pg_base = 0u, pg_mask = 0u, pg_valid = 0u;
    NodeRegs n;
    for (;;)
    {
        if (pg_mask)
        {
            const int bit = find_msb(pg_mask);
            pg_mask &= ~(1u << bit);
            load_record(reinterpret_cast<const char*>(s.prims) + (size_t)prim_record(pg_base, pg_valid, bit) * 48u, 3, n);
            const float v1x = u2f(n.w[0]), v1y = u2f(n.w[1]), v1z = u2f(n.w[2]);
            const float e1x = u2f(n.w[4]), e1y = u2f(n.w[5]), e1z = u2f(n.w[6]);
            const float e2x = u2f(n.w[8]), e2y = u2f(n.w[9]), e2z = u2f(n.w[10]);
            float t = -1.0f, u = 0.0f, v = 0.0f;
            if (wt_hits(w, r.ox, r.oy, r.oz, v1x, v1y, v1z, e1x, e1y, e1z, e2x, e2y, e2z)
                && mt_tuv(r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, v1x, v1y, v1z, e1x, e1y, e1z, e2x, e2y, e2z, t, u, v)
                && t >= 0.0f && t < limit)
            {
                best.t = t; best.u = u; best.v = v; best.hit = 1;
                best.prim_id = n.w[3]; best.geom_id = n.w[7]; best.list_index = n.w[11];
                limit = t;
                if (ANY) return;
            }
            if (pg_mask) continue;
        }
        else
        {
            const int bit = find_msb(ng_mask);
            ng_mask &= ~(1u << bit);
            const uint32_t imask = ng_mask & 0xffu;
            const uint32_t slot = ((uint32_t)(bit - 24) ^ (r.octinv4 & 0xffu)) & 7u;
            const uint32_t rel = (uint32_t)pop_count(imask & ~(0xffffffffu << slot));
            if (ng_mask & 0xff000000u)
            {
                if (sp >= VSB_STACK_SIZE) VSB_STACK_OVERFLOW();
                sb[sp] = ng_base; sm[sp] = ng_mask; ++sp;
            }
            load_record(reinterpret_cast<const char*>(s.nodes + (ng_base + rel)), 5, n);
            uint32_t ph;
            const uint32_t nh = test_node(n, r, limit, ph);
            ng_base = n.w[4]; ng_mask = nh | (n.w[3] >> 24);
            pg_base = n.w[5]; pg_mask = ph; pg_valid = n.w[6];
            if (pg_mask) continue;
        }
        if ((ng_mask & 0xff000000u) == 0u)
        {
            if (sp == 0) return;
            --sp; ng_base = sb[sp]; ng_mask = sm[sp];
        }
    }
}

#if defined(__CUDACC__)
// one thread per ray, rays / hits in the caller's order; over the range of bvhs like traverse_linear.inl:71-130
template <bool ANY, typename SceneTable>
__global__ void __launch_bounds__(128) watertight_kernel(const SceneTable* __restrict__ scenes, int n_scenes, const Ray32* __restrict__ rays,
                                                         unsigned long long n_rays, float max_t, const float* __restrict__ max_t_per_ray, Hit32* hits)
{
    const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rays) return;
    const Ray32 ray = rays[i];
    RayCtx rc;
    ray_ctx_init(rc, ray.ox, ray.oy, ray.oz, ray.dx, ray.dy, ray.dz);
    HitReg best; hit_reset(best);
    const float mt = max_t_per_ray ? max_t_per_ray[i] : max_t;
    float limit = mt;
    for (int k = 0; k < n_scenes; ++k)
    {
        trace_wide_watertight<ANY>(scenes[k].wide, rc, best, limit);
        if (ANY && best.hit) break;
    }
    Hit32 h;
    h.hit = best.hit; h.prim_id = best.prim_id; h.geom_id = best.geom_id; h.t = best.t; h.u = best.u; h.v = best.v;
    h.prim_list_index = best.list_index; h.pad = 0u;
    hits[i] = h;
}
#endif

} // namespace vsb
