// bvh8_traverse.cuh -- ray vs BVH8: the 8-child quantised slab test and the traversal loop.
// Replaces the while-while BVH2 loop of detail/bvh/intersect.inl:35-134 on the fast path.
#pragma once

#include "bvh8.cuh"

namespace vsb
{

struct WideScene
{
    const Node80* nodes;      // BVH8 nodes, root = 0
    const void*   prims;      // TriRec[] or SphRec[] in leaf order
    int           prim_kind;  // 0 triangles, 1 spheres
};

struct RayCtx
{
    float ox, oy, oz, dx, dy, dz;
    float ix, iy, iz;          // 1/dir with |dir| clamped away from 0 (box test only)
    uint32_t octinv4;          // (7 - octant) replicated in 4 bytes
};

VSB_HD void ray_ctx_init(RayCtx& r, float ox, float oy, float oz, float dx, float dy, float dz)
{
    r.ox = ox; r.oy = oy; r.oz = oz; r.dx = dx; r.dy = dy; r.dz = dz;
    const float eps = 8.2718061e-25f;   // 2^-80: keeps (p - o) * (1/eps) finite
    float cx = fabsf(dx) > eps ? dx : (f2u(dx) >> 31 ? -eps : eps);
    float cy = fabsf(dy) > eps ? dy : (f2u(dy) >> 31 ? -eps : eps);
    float cz = fabsf(dz) > eps ? dz : (f2u(dz) >> 31 ? -eps : eps);
    r.ix = 1.0f / cx; r.iy = 1.0f / cy; r.iz = 1.0f / cz;
    uint32_t oct = (cx < 0.0f ? 1u : 0u) | (cy < 0.0f ? 2u : 0u) | (cz < 0.0f ? 4u : 0u);
    r.octinv4 = (7u - oct) * 0x01010101u;
}

struct NodeRegs { uint32_t w[20]; };

VSB_HD void load_node(const Node80* __restrict__ nodes, uint32_t idx, NodeRegs& n)
{
#if defined(__CUDA_ARCH__)
    const uint4* p = reinterpret_cast<const uint4*>(nodes + idx);   // 80 B records, 16-B aligned
    uint4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2), d = __ldg(p + 3), e = __ldg(p + 4);
    n.w[0] = a.x; n.w[1] = a.y; n.w[2] = a.z; n.w[3] = a.w;
    n.w[4] = b.x; n.w[5] = b.y; n.w[6] = b.z; n.w[7] = b.w;
    n.w[8] = c.x; n.w[9] = c.y; n.w[10] = c.z; n.w[11] = c.w;
    n.w[12] = d.x; n.w[13] = d.y; n.w[14] = d.z; n.w[15] = d.w;
    n.w[16] = e.x; n.w[17] = e.y; n.w[18] = e.z; n.w[19] = e.w;
#else
    memcpy(n.w, nodes[idx].w, sizeof(n.w));
#endif
}

// Tests the 8 children of a node against the ray; returns the hit mask: bits 24..31 = inner
// children (bit = 24 + (slot ^ octinv), so the highest set bit is the nearest slot), bits
// 0..23 = leaf primitives (one bit per primitive of every hit leaf child).
VSB_HD uint32_t test_node(const NodeRegs& n, const RayCtx& r, float limit)
{
    const float px = u2f(n.w[0]), py = u2f(n.w[1]), pz = u2f(n.w[2]);
    const uint32_t e = n.w[3];
    // grid scale per axis, pre-multiplied by 2^15 for the byte->float trick: the byte b enters
    // as f = 1 + b * 2^-15, so t = f * S + O with S = s * idir * 2^15, O = (p - o) * idir - S
    const float sxs = u2f(((e & 0xffu) + 15u) << 23) * r.ix;
    const float sys = u2f((((e >> 8) & 0xffu) + 15u) << 23) * r.iy;
    const float szs = u2f((((e >> 16) & 0xffu) + 15u) << 23) * r.iz;
    const float oxs = (px - r.ox) * r.ix;
    const float oys = (py - r.oy) * r.iy;
    const float ozs = (pz - r.oz) * r.iz;
    // error bound of this evaluation + of the reference's own slab arithmetic (see bvh8.cuh):
    // 2^-20 * (|O| + 255 |s idir|) + 2^-8 |s idir|, with |s idir| = |S| * 2^-15
    const float k1 = 9.5367431640625e-07f;                 // 2^-20
    const float k2 = (255.0f * 9.5367431640625e-07f + 0.00390625f) * 3.0517578125e-05f;   // per |S|
    const float mx = fma_f(fabsf(sxs), k2, fabsf(oxs) * k1);
    const float my = fma_f(fabsf(sys), k2, fabsf(oys) * k1);
    const float mz = fma_f(fabsf(szs), k2, fabsf(ozs) * k1);
    const float onx = (oxs - sxs) - mx, ofx = (oxs - sxs) + mx;
    const float ony = (oys - sys) - my, ofy = (oys - sys) + my;
    const float onz = (ozs - szs) - mz, ofz = (ozs - szs) + mz;

    // near/far planes by direction sign (selected once per node, not per child)
    const bool nx = r.ix < 0.0f, ny = r.iy < 0.0f, nz = r.iz < 0.0f;
    uint32_t hitmask = 0;
#pragma unroll
    for (int half = 0; half < 2; ++half)
    {
        const uint32_t meta4 = n.w[6 + half];
        const uint32_t qlox = n.w[8 + half], qloy = n.w[10 + half], qloz = n.w[12 + half];
        const uint32_t qhix = n.w[14 + half], qhiy = n.w[16 + half], qhiz = n.w[18 + half];
        const uint32_t qnx = nx ? qhix : qlox, qfx = nx ? qlox : qhix;
        const uint32_t qny = ny ? qhiy : qloy, qfy = ny ? qloy : qhiy;
        const uint32_t qnz = nz ? qhiz : qloz, qfz = nz ? qloz : qhiz;
        const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
        const uint32_t inner_mask4 = (is_inner4 >> 4) * 0xffu;
        const uint32_t bit_index4 = (meta4 ^ (r.octinv4 & inner_mask4)) & 0x1f1f1f1fu;
        const uint32_t count4 = (meta4 >> 5) & 0x07070707u;
#pragma unroll
        for (int k = 0; k < 4; ++k)
        {
            const float tnx = fma_f(byte_as_unit_float(qnx, k), sxs, onx);
            const float tny = fma_f(byte_as_unit_float(qny, k), sys, ony);
            const float tnz = fma_f(byte_as_unit_float(qnz, k), szs, onz);
            const float tfx = fma_f(byte_as_unit_float(qfx, k), sxs, ofx);
            const float tfy = fma_f(byte_as_unit_float(qfy, k), sys, ofy);
            const float tfz = fma_f(byte_as_unit_float(qfz, k), szs, ofz);
            const float tmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.0f));
            const float tmax = fminf(fminf(tfx, tfy), fminf(tfz, limit));
            const uint32_t cnt = (count4 >> (8 * k)) & 0xffu;
            const uint32_t pos = (bit_index4 >> (8 * k)) & 0xffu;
            if (tmin <= tmax) hitmask |= bit_mask(pos, cnt);   // empty slots have cnt = 0
        }
    }
    return hitmask;
}

// is_closer(hr, result, max_t): update_if.h:50-56,75-79.  `limit` = min(result.t, max_t).
VSB_HD void test_prim(const WideScene& s, uint32_t index, const RayCtx& r, HitReg& best, float& limit)
{
    if (s.prim_kind == 0)
    {
#if defined(__CUDA_ARCH__)
        const float4* p = reinterpret_cast<const float4*>(reinterpret_cast<const TriRec*>(s.prims) + index);
        const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
#else
        const TriRec& tr = reinterpret_cast<const TriRec*>(s.prims)[index];
        struct f4 { float x, y, z, w; };
        const f4 a = { tr.v1x, tr.v1y, tr.v1z, u2f(tr.prim_id) }, b = { tr.e1x, tr.e1y, tr.e1z, u2f(tr.geom_id) },
                 c = { tr.e2x, tr.e2y, tr.e2z, u2f(tr.list_index) };
#endif
        float t, u, v;
        if (isect_triangle(r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, a.x, a.y, a.z, b.x, b.y, b.z, c.x, c.y, c.z, t, u, v)
            && t >= 0.0f && t < limit)
        {
            best.t = t; best.u = u; best.v = v; best.hit = 1;
            best.prim_id = f2u(a.w); best.geom_id = f2u(b.w); best.list_index = f2u(c.w);
            limit = t;
        }
    }
    else
    {
#if defined(__CUDA_ARCH__)
        const float4* p = reinterpret_cast<const float4*>(reinterpret_cast<const SphRec*>(s.prims) + index);
        const float4 a = __ldg(p);
        const uint4 b = __ldg(reinterpret_cast<const uint4*>(p + 1));
#else
        const SphRec& sr = reinterpret_cast<const SphRec*>(s.prims)[index];
        struct f4 { float x, y, z, w; }; struct u4 { uint32_t x, y, z, w; };
        const f4 a = { sr.cx, sr.cy, sr.cz, sr.radius };
        const u4 b = { sr.prim_id, sr.geom_id, sr.list_index, 0u };
#endif
        float t;
        if (isect_sphere(r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, a.x, a.y, a.z, a.w, t) && t >= 0.0f && t < limit)
        {
            best.t = t; best.u = 0.0f; best.v = 0.0f; best.hit = 1;
            best.prim_id = b.x; best.geom_id = b.y; best.list_index = b.z;
            limit = t;
        }
    }
}

#define VSB_STACK_SIZE 48

// Single-ray traversal.  ANY = any-hit (first accepted hit terminates, exit_traversal.h:49-57).
// Stack entries: (base index, mask): node groups carry hit bits in 24..31 and the node's imask
// in 0..7; primitive groups carry one bit per primitive in 0..23.
template <bool ANY>
VSB_HD void trace_wide(const WideScene& s, const RayCtx& r, float max_t, HitReg& best)
{
    hit_reset(best);
    float limit = max_t;
    uint32_t stack_b[VSB_STACK_SIZE], stack_m[VSB_STACK_SIZE];
    int sp = 0;
    uint32_t ng_base = 0u, ng_mask = 0x80000000u;   // root: one "inner child" at bit 31, imask bit 0..7 unused
    // the root is addressed as child_base 0 + popc(imask below slot) = 0: give it imask = 0
    uint32_t pg_base = 0u, pg_mask = 0u;

    for (;;)
    {
        if (ng_mask & 0xff000000u)
        {
            const int bit = find_msb(ng_mask);
            ng_mask &= ~(1u << bit);
            const uint32_t base = ng_base, imask = ng_mask & 0xffu;
            if (ng_mask & 0xff000000u) { stack_b[sp] = ng_base; stack_m[sp] = ng_mask; ++sp; }
            const uint32_t slot = ((uint32_t)(bit - 24) ^ (r.octinv4 & 0xffu)) & 7u;
            const uint32_t rel = (uint32_t)pop_count(imask & ~(0xffffffffu << slot));
            NodeRegs n;
            load_node(s.nodes, base + rel, n);
            const uint32_t hm = test_node(n, r, limit);
            ng_base = n.w[4]; ng_mask = (hm & 0xff000000u) | (n.w[3] >> 24);
            pg_base = n.w[5]; pg_mask = hm & 0x00ffffffu;
        }
        else
        {
            pg_base = ng_base; pg_mask = ng_mask;
            ng_base = 0u; ng_mask = 0u;
        }

        while (pg_mask)
        {
            const int bit = find_msb(pg_mask);
            pg_mask &= ~(1u << bit);
            test_prim(s, pg_base + (uint32_t)bit, r, best, limit);
            if (ANY && best.hit) return;
        }

        if ((ng_mask & 0xff000000u) == 0u)
        {
            if (sp == 0) break;
            --sp;
            ng_base = stack_b[sp]; ng_mask = stack_m[sp];
        }
    }
}

} // namespace vsb
